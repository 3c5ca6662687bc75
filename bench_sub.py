"""Sub-records of bench.py at N = 1 (BASELINE.json configs[2] and configs[4]): a paired-end batch with damaged mates on the
100 Mbp reference (mem_pestat, mate rescue, mem_pair, paired SAM) and long reads on the 3.1 Gbp index (wide-band
ksw_extend2, z-drop, mem_seed_sw), each with its own parity sample against the reference CPU code and its own CPU rate."""
from __future__ import annotations

import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


def _ref_index(pac, meta, bwt, sa, contigs, contig_len):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refapi
    if not refapi.available():
        return None
    rl = refapi.RefLib()
    lens = np.full(contigs, contig_len, dtype=np.int64)
    ann_off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
    return rl.index_from_arrays(np.asarray(bwt), meta["primary"], meta["L2"], meta["seq_len"], np.asarray(sa), meta["sa_intv"],
                                np.asarray(pac), meta["l_pac"], ann_off, lens.astype(np.int32))


def long_reads(a, ctx, hidx, pac, cores, n_reads=100_000, n_check=300):
    """configs[4] shape: reads of 1-10 kbp (3 % sub, 2 % ins, 2 % del, junk tails) on the bench index: worker1 on the device."""
    from bwa_mem_gpu_b200 import abi, bench_index as BI, synth
    meta, pac_c, bwt, sa = BI.load(BI.cache_dir(ROOT, a.contigs, a.contig_len))
    t0 = time.time()
    # reads are drawn from a window of the genome (the simulator wants numpy contigs); the index is the whole genome
    span = min(a.contig_len, 20_000_000)
    contigs = [BI.unpack_pac_numpy(pac_c, c * a.contig_len, c * a.contig_len + span) for c in range(min(a.contigs, 4))]
    seqs, off = synth.simulate_long_reads(contigs, n_reads, seed=1005)
    t_sim = time.time() - t0
    opt = abi.default_opt()
    ctx.batch_upload(seqs, off)
    ctx.batch_run(opt, want_stage_ms=True)
    st = ctx.batch_run(opt, want_stage_ms=True)
    c = ctx.counters()
    rec = {"workload": f"{n_reads} reads of 1-10 kbp ({int(off[-1])} bases) drawn from the first {len(contigs)} contigs, default scoring, "
                       f"{a.contigs}x{a.contig_len / 1e6:.6g} Mbp index", "path": "mem_align1_core batch (reads -> mem_alnreg_t), resident",
           "reads_per_s": n_reads / (st[5] * 1e-3), "bases_per_s": float(off[-1]) / (st[5] * 1e-3), "ms": st[5],
           "stage_ms": dict(zip(["pack", "smem", "seed_sa", "chain", "extend", "total"], [float(x) for x in st])),
           "ksw_extend2_gcups": c["cells"] / (st[4] * 1e-3) / 1e9 if st[4] > 0 else None, "regions": c["regs"], "read_sim_s": t_sim}
    # reads in host memory -> SAM text in host memory (worker2 of long regions: one region per warp, reg2aln_warp_kernel)
    from bwa_mem_gpu_b200 import lib as L
    nf, no = L.numbered_names(b"lr", 0, n_reads)
    L.reads_to_sam([ctx], opt, hidx.names, seqs, off, nf, no, want_off=False)
    t0 = time.perf_counter()
    sam, _, stt = L.reads_to_sam([ctx], opt, hidx.names, seqs, off, nf, no, want_off=False)
    dt = time.perf_counter() - t0
    rec["reads_to_sam"] = {"reads_per_s": n_reads / dt, "seconds": dt, "sam_bytes": int(len(sam)), "worker1_s": stt[1], "worker2_and_text_s": stt[2],
                           "call": "b200_reads_to_sam (pageable host buffers)"}
    idx = _ref_index(pac_c, meta, bwt, sa, a.contigs, a.contig_len)
    if idx is not None:
        m = min(n_check, n_reads)
        sub_off = off[: m + 1]
        t0 = time.time()
        r0, c0 = idx.align(opt, seqs, sub_off, n_threads=cores, cap_per_read=256)
        dt = time.time() - t0
        r1, c1 = ctx.align(opt, seqs, sub_off)
        rec["parity_sample"] = {"reads": int(m), "regions": int(c0.sum()),
                                "identical_to_reference": bool(np.array_equal(c0, c1) and r0.tobytes() == r1.tobytes())}
        o2 = abi.default_opt(); o2.n_threads = cores
        exp = idx.process_seqs(o2, synth.to_ascii(seqs[: int(off[m])]), sub_off, names=[f"lr{i}" for i in range(m)])
        got, _, _ = L.reads_to_sam([ctx], opt, hidx.names, seqs, sub_off, nf, no, want_off=False)
        rec["parity_sample"]["sam_identical_to_reference"] = bool(got.tobytes() == exp)
        rec["cpu_baseline"] = {"value": m / dt, "unit": "reads/s", "cores": cores, "kind": "reference",
                               "sample": f"first {m} reads, mem_align1_core via kt_for, {dt:.2f} s",
                               "e2e_value": m / idx.last_process_seqs_s, "e2e_sample": f"mem_process_seqs (reads -> SAM) {idx.last_process_seqs_s:.2f} s"}
    return rec


def run(a, ctx, hidx, pac, cores):
    out = {}
    try:
        out["long_reads"] = long_reads(a, ctx, hidx, pac, cores)
    except (Exception, SystemExit) as e:  # noqa: BLE001
        out["long_reads"] = {"error": str(e)[:400]}
    ctx.close()                                            # frees the 3.1 Gbp index and the long-read scratch for the next record
    try:
        import bench_pe
        out["paired_end"] = bench_pe.run(a, cores)
    except ImportError:
        out["paired_end"] = {"error": "not built"}
    except (Exception, SystemExit) as e:  # noqa: BLE001
        out["paired_end"] = {"error": str(e)[:400]}
    return out
