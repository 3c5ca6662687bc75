"""Reads in host memory -> SAM text in host memory, single-end (SURVEY.md 8(d) "FASTQ in memory -> SAM text", 8(f) ranks
1-2): b200_align_batch (worker1), b200_reg2sam_se_batch (worker2 before the text: primary marking, MAPQ, record selection,
XA choice, mem_reg2aln) and b200_sam_se_format (mem_aln2sam; host threads) on N reads of the configs[1] workload.  Times the
three host-buffer calls and, on a sample, the reference's worker2 (mem_mark_primary_se + mem_reg2sam, ONE host thread: the
shim loop is serial); checks the sample's SAM text for identity.
usage: python scripts/sam_se_bench.py [--reads N] [--check M] [--threads T]"""
import argparse, json, os, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from bwa_mem_gpu_b200 import abi, lib as L, synth
import refapi


def nt4(seqs):
    lut = np.full(256, 4, dtype=np.uint8)
    for k, ch in enumerate("ACGT"):
        lut[ord(ch)] = lut[ord(ch.lower())] = k
    lut[:4] = np.arange(4)
    return lut[seqs]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=2_000_000)
    ap.add_argument("--check", type=int, default=100_000)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 8)
    a = ap.parse_args()
    contigs = synth.make_genome([25_000_000] * 4)
    reads, _ = synth.simulate_reads(contigs, a.reads, seed=1002)
    seqs = np.ascontiguousarray(reads.reshape(-1)); off = np.arange(0, a.reads + 1, dtype=np.int64) * 150
    names = [f"r{i}" for i in range(a.reads)]
    ctx = L.Context(0)
    hidx = ctx.build_index_device(contigs)
    opt = abi.default_opt()
    t = {}
    for rep in range(2):                                   # the first pass sizes the grow-only buffers
        t0 = time.time()
        regs, cnt = ctx.align(opt, seqs, off)
        t["align_s"] = time.time() - t0
        t0 = time.time()
        regs2, sel, aln, cig, md = ctx.reg2sam_se(opt, seqs, off, regs, cnt)
        t["reg2sam_se_s"] = time.time() - t0
    t0 = time.time()
    sam, soff = ctx.sam_se(opt, hidx.names, names, seqs, off, cnt, sel, aln, cig, md, n_threads=a.threads)
    t["sam_format_s"] = time.time() - t0                   # includes building the ctypes name arrays in Python
    tot = t["align_s"] + t["reg2sam_se_s"] + t["sam_format_s"]
    res = {"reads": a.reads, "regions": int(len(regs)), "records": int((sel["kind"] & 1).sum()), "sam_bytes": len(sam),
           **t, "reads_per_s_reads_to_sam": a.reads / tot, "format_threads": a.threads,
           "mapq_mean": float(sel["mapq"][(sel["kind"] & 1) != 0].mean())}
    if refapi.available():
        m = min(a.check, a.reads)
        rl = refapi.RefLib()
        ridx = rl.index_from_arrays(hidx.bwt_words(), hidx.primary, hidx.L2, hidx.seq_len, hidx.sa(), hidx.sa_intv,
                                    hidx.pac(), hidx.l_pac, hidx.ann_off, hidx.ann_len)
        k = int(cnt[:m].sum())
        t0 = time.time()
        r0, q0, sam0 = ridx.reg2sam_se(opt, nt4(seqs[: 150 * m]), off[: m + 1], regs[:k], cnt[:m])
        t_ref = time.time() - t0
        res.update({"ref_sample_reads": m, "ref_worker2_reads_per_s_1thread": m / t_ref,
                    "regions_identical": bool(r0.tobytes() == regs2[:k].tobytes()),
                    "sam_identical_to_reference": bool(sam0 == sam[: int(soff[m])]), "host_cores": os.cpu_count()})
    print(json.dumps(res))


if __name__ == "__main__":
    main()
