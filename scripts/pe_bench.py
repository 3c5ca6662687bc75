"""Paired-end measurement (BASELINE.json configs[2]): worker1 on the device, mem_pestat on the host (the reference's
function from oracle/_ref, as the host program would), mate rescue on the device.  Prints one JSON line with the
device times, the ksw_align2 call / cell counts (GCUPS) and, on a sample, a parity check of the rescue against the
reference loop plus the reference's CPU time for it.  Measurement helper, not the product path.

usage: python scripts/pe_bench.py [--pairs N] [--damaged F] [--contigs 4 --contig-len 25000000] [--check-pairs M]
"""
import argparse, json, os, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from bwa_mem_gpu_b200 import abi, lib as L, synth
import refapi


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=1_000_000)
    ap.add_argument("--damaged", type=float, default=0.05)
    ap.add_argument("--contigs", type=int, default=4)
    ap.add_argument("--contig-len", type=int, default=25_000_000)
    ap.add_argument("--check-pairs", type=int, default=20000)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    contigs = synth.make_genome([a.contig_len] * a.contigs)
    reads, _ = synth.simulate_pairs(contigs, a.pairs, seed=77, damaged_frac=a.damaged, damage="periodic")
    seqs = np.ascontiguousarray(reads.reshape(-1))
    n = 2 * a.pairs
    off = np.arange(0, n + 1, dtype=np.int64) * reads.shape[1]
    t0 = time.time()
    hidx = L.HostIndex(contigs)
    ctx = L.Context(0)
    ctx.upload_host_index(hidx)
    t_index = time.time() - t0
    opt = abi.default_opt(); opt.flag |= abi.MEM_F_PE
    rl = refapi.RefLib()
    ridx = rl.index_from_arrays(hidx.bwt_words(), hidx.primary, hidx.L2, hidx.seq_len, hidx.sa(), hidx.sa_intv,
                                hidx.pac(), hidx.l_pac, hidx.ann_off, hidx.ann_len)
    ctx.batch_upload(seqs, off)
    res = {}
    for rep in range(a.reps):
        ms = ctx.batch_run(opt, want_stage_ms=True)
        regs, cnt = ctx.batch_download(n, 3 * n)
        if rep == 0:
            pes = ridx.pestat(opt, regs, cnt)
        n_sw, cells, rms = ctx.batch_rescue(opt, pes)
        res = {"worker1_ms": ms[5], "rescue_ms": rms, "ksw_align2_calls": n_sw, "cells": cells,
               "gcups": cells / rms / 1e6 if rms else 0.0, "pairs_per_s_device": a.pairs / ((ms[5] + rms) * 1e-3)}
    g_regs, g_cnt = ctx.batch_download(n, 3 * n)
    # parity + CPU time of the reference loop on the first check-pairs pairs (same pes)
    m = 2 * min(a.check_pairs, a.pairs)
    k = int(cnt[:m].sum())
    t0 = time.time()
    e_regs, e_cnt, e_sw = ridx.mate_rescue(opt, pes, seqs, off[: m + 1], regs[:k], cnt[:m])
    t_ref = time.time() - t0
    kk = int(g_cnt[:m].sum())
    ok = bool(np.array_equal(e_cnt, g_cnt[:m]) and e_regs.tobytes() == g_regs[:kk].tobytes())
    res.update({"pairs": a.pairs, "damaged_frac": a.damaged, "index_s": round(t_index, 1),
                "pes_FR": [int(pes["low"][1]), int(pes["high"][1])], "parity_sample_pairs": m // 2, "parity": ok,
                "ref_rescue_cpu_s_1thread": round(t_ref, 3), "ref_sw_calls_sample": int(e_sw)})
    print(json.dumps(res))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
