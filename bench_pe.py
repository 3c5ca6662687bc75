"""Paired-end sub-record of bench.py (BASELINE.json configs[2]: 100 Mbp reference, 2 x 150 bp FR pairs, insert N(500, 50),
5 % of the pairs with one damaged mate so that mem_matesw / ksw_align2 runs): pairs in pinned host memory -> paired SAM
text in pinned host memory through b200_reads_to_sam with MEM_F_PE (worker1, mem_pestat, mate rescue, mem_pair, SAM on the
device), a parity sample against the reference's mem_process_seqs and the reference's rate on the host cores."""
from __future__ import annotations

import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
CONTIGS, CONTIG_LEN = 4, 25_000_000


def pair_names(L, n_pairs):
    flat, off = L.numbered_names(b"p", 0, n_pairs)
    lens = np.diff(off)
    lens2 = np.repeat(lens, 2)
    off2 = np.zeros(2 * n_pairs + 1, dtype=np.int64)
    np.cumsum(lens2, out=off2[1:])
    src0 = np.repeat(off[:-1], 2)
    pos = np.arange(int(off2[-1]), dtype=np.int64) - np.repeat(off2[:-1], lens2) + np.repeat(src0, lens2)
    return np.ascontiguousarray(flat[pos]), off2


def run(a, cores, n_pairs=1_000_000, n_check=20_000, cpu_pairs=150_000, reps=3, device=0):
    import torch
    from bwa_mem_gpu_b200 import abi, bench_index as BI, lib as L, synth
    import argparse
    cfg = argparse.Namespace(contigs=CONTIGS, contig_len=CONTIG_LEN)
    sys.path.insert(0, ROOT)
    import bench
    d = bench.ensure_index_cache(cfg, device)
    meta, pac, bwt, sa = BI.load(d)
    contigs = [BI.unpack_pac_numpy(pac, c * CONTIG_LEN, (c + 1) * CONTIG_LEN) for c in range(CONTIGS)]
    t0 = time.time()
    reads, _ = synth.simulate_pairs(contigs, n_pairs, seed=1003, damaged_frac=0.05)
    t_sim = time.time() - t0
    n = 2 * n_pairs
    pin = lambda nb: torch.empty(int(nb), dtype=torch.uint8).pin_memory().numpy()
    seqs = pin(n * 150); seqs[:] = reads.reshape(-1); del reads
    off = pin((n + 1) * 8).view(np.int64); off[:] = np.arange(0, n + 1, dtype=np.int64) * 150
    nf0, no0 = pair_names(L, n_pairs)
    nf = pin(len(nf0) + 1); nf[: len(nf0)] = nf0
    no = pin(no0.nbytes).view(np.int64); no[:] = no0
    ctx = L.Context(device)
    hidx = ctx.build_index_device_pac(np.asarray(pac), [CONTIG_LEN] * CONTIGS)
    opt = abi.default_opt()
    opt.flag |= abi.MEM_F_PE
    sam = pin(n * 460 + 4096)
    times, st = [], None
    for rep in range(reps + 1):
        t0 = time.perf_counter()
        out, _, st = L.reads_to_sam([ctx], opt, hidx.names, seqs, off, nf, no, sam=sam, want_off=False)
        if rep:
            times.append(time.perf_counter() - t0)
    dt = float(np.mean(times))
    rec = {"workload": f"{n_pairs} pairs 2x150 bp (FR, insert N(500,50), 1% sub, 5% of pairs with one mate damaged) on {CONTIGS}x{CONTIG_LEN / 1e6:g} Mbp, "
                       "one call = one mem_pestat batch",
           "path": "b200_reads_to_sam with MEM_F_PE: pairs in pinned host memory -> paired SAM text in pinned host memory",
           "pairs_per_s": n_pairs / dt, "reads_per_s": n / dt, "seconds": dt, "sam_bytes": int(len(out)),
           "thread_seconds": {"worker1_rescue": st[1], "worker2_and_text": st[2]}, "read_sim_s": t_sim,
           "proper_pair_records": int(np.count_nonzero(np.frombuffer(out[: min(len(out), 50_000_000)].tobytes(), dtype=np.uint8) == 10))}
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refapi
    if refapi.available():
        rl = refapi.RefLib()
        lens = np.full(CONTIGS, CONTIG_LEN, dtype=np.int64)
        ann_off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
        idx = rl.index_from_arrays(np.asarray(bwt), meta["primary"], meta["L2"], meta["seq_len"], np.asarray(sa), meta["sa_intv"],
                                   np.asarray(pac), meta["l_pac"], ann_off, lens.astype(np.int32))
        o2 = abi.default_opt(); o2.flag |= abi.MEM_F_PE; o2.n_threads = cores
        m = 2 * min(n_check, n_pairs)                       # the sample as a batch of its own on both sides (same mem_pestat batch)
        names = [f"p{i // 2}" for i in range(m)]
        sam0 = idx.process_seqs(o2, synth.to_ascii(seqs[: m * 150]), off[: m + 1], names=names, n_processed=0)
        got, _, _ = L.reads_to_sam([ctx], opt, hidx.names, seqs, off[: m + 1], nf, no, want_off=False)
        rec["parity_sample"] = {"pairs": m // 2, "sam_bytes": len(sam0), "sam_identical_to_reference": bool(got.tobytes() == sam0),
                                "records_with_mate_rescue_tag_free_check": None}
        mc = 2 * min(cpu_pairs, n_pairs)
        names = [f"p{i // 2}" for i in range(mc)]
        idx.process_seqs(o2, synth.to_ascii(seqs[: mc * 150]), off[: mc + 1], names=names, n_processed=0)
        rec["cpu_baseline"] = {"value": mc / idx.last_process_seqs_s, "unit": "reads/s", "cores": cores, "kind": "reference",
                               "sample": f"first {mc // 2} pairs, mem_process_seqs with MEM_F_PE (reads -> SAM), {idx.last_process_seqs_s:.2f} s"}
        rec["parity_sample"].pop("records_with_mate_rescue_tag_free_check")
    ctx.close()
    return rec


if __name__ == "__main__":
    import argparse, json
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=1_000_000)
    a = ap.parse_args()
    print(json.dumps(run(a, os.cpu_count() or 1, n_pairs=a.pairs)))
