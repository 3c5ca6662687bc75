"""A/B of context tunables on the long-read workload (configs[4] shape) over the cached bench index.
    python scripts/long_ab.py [--small] --reads 50000 --set smem_path=3 --set smem_lcap=24"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from bwa_mem_gpu_b200 import abi, bench_index as BI, lib as L, synth
ap = argparse.ArgumentParser()
ap.add_argument("--small", action="store_true")
ap.add_argument("--reads", type=int, default=50000)
ap.add_argument("--set", action="append", default=[])
ap.add_argument("--sam", action="store_true", help="also time reads -> SAM text (b200_reads_to_sam) and check a sample against the reference")
a = ap.parse_args()
a.contigs, a.contig_len = (4, 25_000_000) if a.small else (24, 129_166_667)
d = bench.ensure_index_cache(a)
meta, pac, bwt, sa = BI.load(d)
span = min(a.contig_len, 20_000_000)
contigs = [BI.unpack_pac_numpy(pac, c * a.contig_len, c * a.contig_len + span) for c in range(4)]
seqs, off = synth.simulate_long_reads(contigs, a.reads, seed=1005)
opt = abi.default_opt()
base = None
for s in a.set or [""]:
    ctx = L.Context(0)
    for kv in filter(None, s.split(",")):
        k, v = kv.split("="); ctx.tune(k, int(v))
    ctx.build_index_device_pac(np.asarray(pac), [a.contig_len] * a.contigs)
    ctx.batch_upload(seqs, off)
    ctx.batch_run(opt, want_stage_ms=True)
    st = ctx.batch_run(opt, want_stage_ms=True)
    c = ctx.counters()
    base = base or c["regs"]
    extra = {}
    if a.sam:
        import time
        nf, no = L.numbered_names(b"lr", 0, a.reads)
        hn = [f"chr{i + 1}" for i in range(a.contigs)]
        L.reads_to_sam([ctx], opt, hn, seqs, off, nf, no, want_off=False)
        t0 = time.perf_counter()
        sam, _, stt = L.reads_to_sam([ctx], opt, hn, seqs, off, nf, no, want_off=False)
        dt = time.perf_counter() - t0
        extra = {"reads_to_sam_s": dt, "reads_to_sam_reads_per_s": a.reads / dt, "sam_bytes": len(sam), "worker1_s": stt[1], "worker2_text_s": stt[2]}
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import refapi
        if refapi.available() and not extra.get("checked"):
            rl = refapi.RefLib()
            lens = np.full(a.contigs, a.contig_len, dtype=np.int64)
            ann_off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
            idx = rl.index_from_arrays(np.asarray(bwt), meta["primary"], meta["L2"], meta["seq_len"], np.asarray(sa), meta["sa_intv"], np.asarray(pac), meta["l_pac"], ann_off, lens.astype(np.int32))
            m = min(300, a.reads)
            o2 = abi.default_opt(); o2.n_threads = os.cpu_count() or 1
            exp = idx.process_seqs(o2, synth.to_ascii(seqs[: int(off[m])]), off[: m + 1], names=[f"lr{i}" for i in range(m)])
            extra["ref_process_seqs_reads_per_s"] = m / idx.last_process_seqs_s
            got, _, _ = L.reads_to_sam([ctx], opt, hn, seqs, off[: m + 1], nf, no, want_off=False)
            extra["sam_sample_identical"] = bool(got.tobytes() == exp)
    print(json.dumps({"set": s, "reads": a.reads, **extra, "bases": int(off[-1]), "reads_per_s": a.reads / (st[5] * 1e-3),
                      "stage_ms": dict(zip(["pack", "smem", "seed_sa", "chain", "extend", "total"], [round(float(x), 1) for x in st])),
                      "regs": c["regs"], "same_regs": c["regs"] == base}), flush=True)
    ctx.close()
