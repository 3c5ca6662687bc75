"""A/B of seeding tunables on ONE genome (default: the 3.1 Gbp configs[3] reference): for every --set "k=v,k=v" a fresh
context builds the index on the device, runs the resident batch 1 + reps times and prints stage times + counters.
    python scripts/seed_ab.py --human --reads 4000000 --set kmer_k=9 --set kmer_k=12 --set kmer_k=13"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bwa_mem_gpu_b200 import abi, lib as L, synth

ap = argparse.ArgumentParser()
ap.add_argument("--human", action="store_true")
ap.add_argument("--contigs", type=int, default=4)
ap.add_argument("--contig-len", type=int, default=25_000_000)
ap.add_argument("--reads", type=int, default=4_000_000)
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--set", action="append", default=[])
ap.add_argument("--lib", default=None, help="alternative build of the product library")
a = ap.parse_args()
if a.human:
    a.contigs, a.contig_len = 24, 129_166_667
contigs = synth.make_genome([a.contig_len] * a.contigs)
reads, _ = synth.simulate_reads(contigs, a.reads, seed=1002)
seqs = np.ascontiguousarray(reads.reshape(-1)); off = np.arange(0, a.reads + 1, dtype=np.int64) * 150
opt = abi.default_opt()
base = None
for s in a.set or [""]:
    ctx = L.Context(0, L.load(a.lib) if a.lib else None)
    for kv in filter(None, s.split(",")):
        k, v = kv.split("="); ctx.tune(k, int(v))
    t0 = time.time(); ctx.build_index_device(contigs); t_idx = time.time() - t0
    ctx.batch_upload(seqs, off)
    ctx.batch_run(opt, want_stage_ms=True)
    st = np.zeros(6)
    for _ in range(a.reps):
        st += np.array(ctx.batch_run(opt, want_stage_ms=True))
    st /= a.reps
    ctx.tune("seed_stats", 1); ctx.batch_run(opt, want_stage_ms=True)      # counters come from one extra, untimed run
    c = ctx.counters()
    if base is None:
        base = c["regs"]
    print(json.dumps({"set": s, "index_s": round(t_idx, 2), "reads": a.reads, "smem_ms": st[1], "total_ms": st[5],
                      "smem_ms_per_2^20": st[1] * (1 << 20) / a.reads, "Mreads_s": a.reads / st[5] / 1e3,
                      "stage_ms": [round(float(x), 2) for x in st], "extend": c["extend"], "two_block": c["two_block"],
                      "regs": c["regs"], "regs_same_as_first": c["regs"] == base}), flush=True)
    ctx.close()
