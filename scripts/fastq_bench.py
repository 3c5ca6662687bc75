"""Input side measured: a FASTQ file of N reads written on the box, then (a) b200_fastq_read alone (plain and gzip),
(b) FASTQ -> SAM through b200_fastq_to_sam (reader | devices | writer), next to (c) the in-memory b200_reads_to_sam rate on the
same reads.  usage: python scripts/fastq_bench.py [--reads N] [--small] [--threads T]"""
import argparse, gzip, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from bwa_mem_gpu_b200 import abi, bench_index as BI, lib as L, synth

ap = argparse.ArgumentParser()
ap.add_argument("--reads", type=int, default=8_000_000)
ap.add_argument("--small", action="store_true")
ap.add_argument("--threads", type=int, default=os.cpu_count() or 8)
ap.add_argument("--chunk", type=int, default=150 * (1 << 20))
a = ap.parse_args()
a.contigs, a.contig_len = (4, 25_000_000) if a.small else (24, 129_166_667)
d = bench.ensure_index_cache(a)
meta, pac, _, _ = BI.load(d)
g = BI.unpack_pac_numpy(pac, 0, min(meta["l_pac"], 4 * a.contig_len))
nc = min(a.contigs, 4)
parts = [synth.counter_reads(np, g, nc, a.contig_len, b, min(a.reads, b + 500_000))[0] for b in range(0, a.reads, 500_000)]
reads = np.concatenate(parts); del parts
n = a.reads
# FASTQ text, vectorised: "@r<i>\n<seq>\n+\n<qual>\n"
nf, no = L.numbered_names(b"r", 0, n)
asc = synth.to_ascii(reads)
rec_len = (no[1:] - no[:-1]) + 1 + 1 + 150 + 1 + 2 + 150 + 1
off = np.zeros(n + 1, dtype=np.int64); np.cumsum(rec_len, out=off[1:])
txt = np.full(int(off[-1]), 10, dtype=np.uint8)
txt[off[:-1]] = ord("@")
name_len = no[1:] - no[:-1]
pos = np.arange(int(no[-1]), dtype=np.int64)
txt[np.repeat(off[:-1] + 1 - no[:-1], name_len) + pos] = nf[: int(no[-1])]
s0 = off[:-1] + 1 + name_len + 1
ar = np.arange(150, dtype=np.int64)
for b in range(0, n, 1 << 20):
    e = min(n, b + (1 << 20))
    txt[(s0[b:e, None] + ar[None, :]).reshape(-1)] = asc[b:e].reshape(-1)
    txt[s0[b:e] + 151] = ord("+")
    txt[(s0[b:e, None] + 153 + ar[None, :]).reshape(-1)] = ord("I")
os.makedirs(os.path.join(ROOT, "tests", "_data"), exist_ok=True)
fq = os.path.join(ROOT, "tests", "_data", "bench_reads.fq")
txt.tofile(fq)
res = {"reads": n, "fastq_bytes": int(len(txt)), "threads": a.threads}
def read_all(path, threads):
    r = L.FastqReader(path, n_threads=threads)
    t0 = time.perf_counter(); tot = 0
    while True:
        b = r.read(a.chunk)
        if b is None: break
        tot += b["n"]
    dt = time.perf_counter() - t0
    r.close()
    assert tot == n
    return dt
for th in (1, a.threads):
    dt = read_all(fq, th)
    res[f"reader_plain_{th}thr"] = {"seconds": dt, "Mreads_s": n / dt / 1e6, "GB_s": len(txt) / dt / 1e9}
ctx = L.Context(0)
hidx = ctx.build_index_device_pac(np.asarray(pac), [a.contig_len] * a.contigs)
ctx2 = L.Context(0); ctx2.replicate_index_from(ctx)
opt = abi.default_opt()
for rep in range(2):
    nr, st = L.fastq_to_sam([ctx, ctx2], opt, hidx.names, fq, chunk_bases=a.chunk, out_path=None, n_threads=a.threads)
res["fastq_to_sam"] = {"seconds": st[0], "Mreads_s": n / st[0] / 1e6, "read_parse_s": st[1], "align_s": st[2], "write_s": st[3], "batches": st[4], "sam_bytes": st[5],
                       "call": "b200_fastq_to_sam, 2 contexts on one GPU, SAM discarded (out_fd < 0)"}
seqs = np.ascontiguousarray(reads.reshape(-1)); o = np.arange(n + 1, dtype=np.int64) * 150
sam = np.zeros(n * 420 + 4096, dtype=np.uint8)
for rep in range(2):
    t0 = time.perf_counter(); L.reads_to_sam([ctx, ctx2], opt, hidx.names, seqs, o, nf, no, sam=sam, want_off=False); dt = time.perf_counter() - t0
res["in_memory_reads_to_sam"] = {"seconds": dt, "Mreads_s": n / dt / 1e6, "note": "pageable host buffers"}
res["fastq_over_in_memory"] = res["fastq_to_sam"]["Mreads_s"] / res["in_memory_reads_to_sam"]["Mreads_s"]
os.remove(fq)
print(json.dumps(res))
