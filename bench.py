#!/usr/bin/env python
"""bench.py -- headline benchmark: aligned reads/s of the BWA-MEM alignment core on B200.

Workload (BASELINE.json configs[1]): synthetic 100 Mbp reference (4 x 25 Mbp, seed 20261104), 10 M x 150 bp
single-end reads (1 % substitutions, 0.015 %/base 1-bp indels) per GPU.  One "step" = one pass of the hot path
(SMEM seeding -> SA location -> chaining/filtering -> banded extension -> dedup) over that read batch,
i.e. worker1 / mem_align1_core of mem_process_seqs for every read: reads in, mem_alnreg_t records out.

  value : whole-job reads/s with the reads already resident in HBM (device pipeline only)
  e2e   : the same through the C ABI call a host program makes (b200_align_batch) with PINNED HOST buffers:
          H2D of the reads and D2H of the mem_alnreg_t records are inside the timed region
  --impl reference : the reference's own CPU code (oracle/_ref, mem_align1_core over kt_for with all host
          threads) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "aligned_reads_per_sec"
UNIT = "reads/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--reads", type=int, default=10_000_000, help="reads per GPU per step")
    ap.add_argument("--contigs", type=int, default=4)
    ap.add_argument("--contig-len", type=int, default=25_000_000)
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("--cpu-sample", type=int, default=0, help="reads in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--tune", default="", help="comma-separated key=value context tunables (b200_ctx_tune)")
    ap.add_argument("--index", default="auto", choices=["auto", "host", "device"],
                    help="index construction: host SA-IS builder or the device builder (b200_index_build_device); auto = device")
    ap.add_argument("--human", action="store_true",
                    help="BASELINE.json configs[3] reference: 24 contigs totalling 3.1 Gbp (device-built index; 64-bit positions)")
    ap.add_argument("--check", type=int, default=20000,
                    help="reads of the step's batch whose mem_alnreg_t records are compared with the reference CPU code (0 = none)")
    a = ap.parse_args()
    if a.human:
        a.contigs, a.contig_len = 24, 129_166_667
    return a


def workload_name(a):
    return (f"synthetic {a.contigs}x{a.contig_len / 1e6:.6g} Mbp reference (seed 20261104), "
            f"{a.reads / 1e6:g}M x {a.read_len} bp single-end reads per GPU, 1% sub + 0.015%/bp indel")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index: int):
        super().__init__(daemon=True)
        self.idx, self.rows, self.stop_flag, self.proc = gpu_index, [], False, None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def build_inputs(a, rank):
    from bwa_mem_gpu_b200 import synth
    contigs = synth.make_genome([a.contig_len] * a.contigs)
    t0 = time.time()
    reads, truth = synth.simulate_reads(contigs, a.reads, read_len=a.read_len, seed=1002 + 7919 * rank)
    seqs = np.ascontiguousarray(reads.reshape(-1))
    off = np.arange(0, a.reads + 1, dtype=np.int64) * a.read_len
    build_inputs.truth = truth
    return contigs, seqs, off, time.time() - t0


def truth_check(a, contigs, truth, regs, cnt):
    """Size-independent property: a read drawn from a known place must come back there.  Fraction of reads whose
    best region (first record) lies on the true strand within 20 bp of the true start."""
    lens = np.array([len(c) for c in contigs], dtype=np.int64)
    l_pac = int(lens.sum())
    starts_c = np.concatenate([[0], np.cumsum(lens)])[:-1]
    n = len(cnt)
    first = np.concatenate([[0], np.cumsum(cnt.astype(np.int64))])[:-1]
    has = cnt > 0
    r = regs[first[has]]
    gpos = (starts_c[truth["contig"]] + truth["pos"])[:n][has]
    rev = truth["rev"][:n][has]
    on_rev = r["rb"] >= l_pac
    fwd_start = np.where(on_rev, 2 * l_pac - r["re"], r["rb"])                  # forward-strand start of the aligned part
    clip = np.where(on_rev, a.read_len - r["qe"], r["qb"])                       # read bases before it on the forward strand
    ok = (on_rev == rev) & (np.abs(fwd_start - clip - gpos) <= 20)
    return {"reads": int(n), "with_region": int(has.sum()), "best_region_at_true_locus": int(ok.sum()),
            "frac": float(ok.sum()) / max(1, n)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


SEED_KERNELS = "smem_p3_kernel+smem_fast3_kernel"


def measured_traffic(kernel):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (profiles/traffic.json:
    dram__bytes_read.sum + dram__bytes_write.sum, one launch = one 2^20-read chunk), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    tab = json.load(open(p))
    tot = 0.0
    for k in kernel.split("+"):                              # a stage of several kernels: the sum of their captures
        if k not in tab:
            return None
        tot += float(tab[k]["dram_bytes_per_launch"])
    return tot


def microbench_ceilings():
    """Random 32-B-sector gather rates and the DPX issue rate measured on this pool (profiles/microbench.json)."""
    p = os.path.join(ROOT, "profiles", "microbench.json")
    return json.load(open(p)) if os.path.exists(p) else None


def reads_to_sam(a, ctx, hidx, opt, pin_seq, pin_off, threads):
    """SURVEY.md 8(d) "reads in memory -> SAM text": the first m reads of the step's batch through b200_align_batch (region
    counts only come back: cap 0), b200_batch_reg2sam_se (worker2 on the resident batch: primary marking, MAPQ, record
    selection, XA, CIGAR/NM/MD) and b200_sam_se_format (host threads); one timed region around the three C calls, pinned
    host buffers for the reads and the records, a pageable buffer for the SAM text."""
    import ctypes as C
    from bwa_mem_gpu_b200 import abi
    m = min(a.reads, 4_000_000)
    P = lambda x, ty: x.ctypes.data_as(C.POINTER(ty))
    seqs = pin_seq.numpy(); off = pin_off.numpy()[: m + 1]
    cnt = np.zeros(m, dtype=np.int32)
    a_names = (C.c_char_p * len(hidx.names))(*[x.encode() for x in hidx.names])
    r_names = (C.c_char_p * m)(*[b"r%d" % i for i in range(m)])
    soff = np.zeros(m + 1, dtype=np.int64)
    sam = np.zeros((a.read_len + 260) * m + 4096, dtype=np.uint8)
    nr0 = -ctx.lib.b200_align_batch(ctx.h, C.byref(opt), C.c_int(m), P(seqs, C.c_uint8), P(off, C.c_int64), C.c_int64(0), None,
                                    P(cnt, C.c_int32))                 # sizes the record buffers (regions per read vary with the genome)
    cap_r = int(max(nr0, m) * 1.02) + 1024
    import torch

    def pinned(n, dt):                                     # pinned host buffers, as a host program feeding a GPU would use
        return torch.zeros(n * np.dtype(dt).itemsize, dtype=torch.uint8).pin_memory().numpy().view(dt)
    sel = pinned(cap_r, abi.SEL_DTYPE); aln = pinned(cap_r, abi.ALN_DTYPE)
    regs2 = np.zeros(cap_r, dtype=abi.ALNREG_DTYPE)
    cig = pinned(4 * cap_r + 64, np.uint32); md = pinned(16 * cap_r + 64, np.uint8)
    tot2 = (C.c_int64 * 2)()
    best = None
    for it in range(3):                                    # the first pass sizes the context's grow-only buffers
        t0 = time.perf_counter()
        nr = -ctx.lib.b200_align_batch(ctx.h, C.byref(opt), C.c_int(m), P(seqs, C.c_uint8), P(off, C.c_int64), C.c_int64(0), None,
                                       P(cnt, C.c_int32))
        t1 = time.perf_counter()
        if nr <= 0 or nr > cap_r:
            return {"error": f"align: {nr}"}, None
        rc = ctx.lib.b200_batch_reg2sam_se(ctx.h, C.byref(opt), C.c_int64(0), P(cnt, C.c_int32),
                                           C.c_void_p(regs2.ctypes.data) if it == 0 else None, C.c_void_p(sel.ctypes.data),
                                           C.c_void_p(aln.ctypes.data), C.c_int64(len(cig)), P(cig, C.c_uint32), C.c_int64(len(md)),
                                           C.c_void_p(md.ctypes.data), tot2)
        t2 = time.perf_counter()
        if rc != nr:
            return {"error": ctx.lib.b200_last_error().decode()}, None
        nb = ctx.lib.b200_sam_se_format(C.byref(opt), a_names, None, C.c_int(m), r_names, None, P(seqs, C.c_uint8), None,
                                        P(off, C.c_int64), P(cnt, C.c_int32), C.c_void_p(sel.ctypes.data), C.c_void_p(aln.ctypes.data),
                                        P(cig, C.c_uint32), C.c_void_p(md.ctypes.data), None, C.c_int(threads), C.c_int64(len(sam)),
                                        C.c_void_p(sam.ctypes.data), P(soff, C.c_int64))
        t3 = time.perf_counter()
        if nb < 0:
            return {"error": f"sam buffer: {nb}"}, None
        if it and (best is None or t3 - t0 < best[0]):
            best = (t3 - t0, t1 - t0, t2 - t1, t3 - t2)
    staged = {"reads_per_s": m / best[0], "align_s": best[1], "worker2_device_s": best[2], "sam_format_s": best[3],
              "path": "b200_align_batch (counts only) -> b200_batch_reg2sam_se (resident batch) -> b200_sam_se_format, one after the other"}
    # the same as ONE call that writes the text of sub-batch k while sub-batch k+1 is on the device
    sam1 = sam.copy(); soff1 = soff.copy()
    t_pipe = None
    for it in range(3):
        t0 = time.perf_counter()
        nb2 = ctx.lib.b200_align_sam_se_batch(ctx.h, C.byref(opt), a_names, None, C.c_int(m), C.c_int64(0), r_names, None,
                                              P(seqs, C.c_uint8), None, P(off, C.c_int64), None, C.c_int(threads), C.c_int(0),
                                              C.c_int64(len(sam)), C.c_void_p(sam.ctypes.data), P(soff, C.c_int64))
        dt = time.perf_counter() - t0
        if nb2 < 0:
            return {"error": f"b200_align_sam_se_batch: {nb2} {ctx.lib.b200_last_error().decode()}"}, None
        if it and (t_pipe is None or dt < t_pipe):
            t_pipe = dt
    same = bool(nb2 == nb and np.array_equal(sam[:nb2], sam1[:nb]) and np.array_equal(soff, soff1))
    out = {"value": m / t_pipe, "unit": UNIT, "reads": m, "regions": int(nr), "sam_bytes": int(nb2), "seconds": t_pipe,
           "format_threads": threads, "pipelined_equals_staged": same,
           "path": "b200_align_sam_se_batch (reads in pinned host memory -> SAM text in host memory, sub-batches of 2^20 reads, "
                   "host formatting overlapped with the device); best of 2 timed passes",
           "staged": staged}
    return out, (cnt, regs2, sam, soff)


def cpu_reference(a, hidx, seqs, off, n_sample, threads, check=None, sam_check=None):
    """The reference's CPU implementation of the same path (mem_align1_core over kt_for), oracle/_ref."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refapi
    from bwa_mem_gpu_b200 import abi
    if not refapi.available():
        return None
    rl = refapi.RefLib()
    idx = rl.index_from_arrays(hidx.bwt_words(), hidx.primary, hidx.L2, hidx.seq_len, hidx.sa(), hidx.sa_intv,
                               hidx.pac(), hidx.l_pac, hidx.ann_off, hidx.ann_len)
    opt = abi.default_opt()
    sub_off = off[: n_sample + 1]
    t0 = time.time()
    _, cnt = idx.align(opt, seqs, sub_off, n_threads=threads, count_only=True)
    dt = time.time() - t0
    out = {"value": n_sample / dt, "unit": UNIT, "cores": threads, "kind": "reference",
           "sample": f"first {n_sample} reads of the step's batch, mem_align1_core via kt_for, {dt:.2f} s",
           "regions": int(cnt.sum())}
    if check is not None:                                  # parity on a sample: all 88 bytes of every record
        n_chk, dev_regs, dev_cnt = check
        r0, c0 = idx.align(opt, seqs, off[: n_chk + 1], n_threads=threads, cap_per_read=64)
        same = bool(np.array_equal(c0, dev_cnt[:n_chk])) and r0.tobytes() == dev_regs[: int(c0.sum())].tobytes()
        out["parity_sample"] = {"reads": int(n_chk), "regions": int(c0.sum()), "identical_to_reference": same}
    if sam_check is not None and check is not None:        # the reference's worker2 (mem_mark_primary_se + mem_reg2sam) on the sample
        cnt_s, regs2, sam, soff = sam_check
        n_chk = min(check[0], len(cnt_s))
        k = int(cnt_s[:n_chk].astype(np.int64).sum())
        regs0 = check[1]                                   # worker1's regions of the sample, as downloaded by the e2e call
        lut = np.full(256, 4, dtype=np.uint8); lut[:4] = np.arange(4)
        for c, ch in enumerate("ACGT"):
            lut[ord(ch)] = lut[ord(ch.lower())] = c
        t0 = time.time()
        r0, _, sam0 = idx.reg2sam_se(opt, lut[seqs[: int(off[n_chk])]], off[: n_chk + 1], regs0[:k], cnt_s[:n_chk])
        dt = time.time() - t0
        out["sam_parity_sample"] = {"reads": int(n_chk), "sam_bytes": len(sam0), "ref_worker2_reads_per_s_1thread": n_chk / dt,
                                    "regions_identical": bool(r0.tobytes() == regs2[:k].tobytes()),
                                    "sam_identical_to_reference": bool(sam0 == sam[: int(soff[n_chk])].tobytes())}
    return out


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = os.cpu_count() or 1

    if a.impl == "reference":
        if rank != 0:
            return
        from bwa_mem_gpu_b200 import lib as L
        n_sample = a.cpu_sample or min(a.reads, max(20000, 150000 * cores))
        contigs, seqs, off, _ = build_inputs(a, 0)
        hidx = L.HostIndex(contigs)          # host-side index construction only (no device work)
        vals = []
        for s in range(a.warmup + a.steps):
            r = cpu_reference(a, hidx, seqs, off, n_sample, cores)
            if r is None:
                print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libbwaref.so not built"}))
                return
            if s >= a.warmup:
                vals.append(r)
        v = float(np.mean([x["value"] for x in vals]))
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * n_sample / v, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": workload_name(a), "path": "mem_align1_core batch (reads -> mem_alnreg_t)",
                       "sample_reads_per_step": n_sample},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": vals[-1]["sample"]},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }))
        return

    import torch
    import torch.distributed as dist
    from bwa_mem_gpu_b200 import abi, lib as L

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU implementation")
    torch.cuda.set_device(local_rank)
    if world > 1:
        # keep stdout to the one JSON line: NCCL writes its version banner to fd 1 when the communicator is created (eager
        # here because of device_id, exercised once by the barrier), so fd 1 points at stderr while that happens
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    t_setup = time.time()
    contigs, seqs, off, t_reads = build_inputs(a, rank)
    ctx = L.Context(local_rank)
    for kv in filter(None, a.tune.split(",")):
        k, v = kv.split("=")
        ctx.tune(k, int(v))
    t0 = time.time()
    if a.index == "host":
        hidx = L.HostIndex(contigs)
        t_index = time.time() - t0
        t0 = time.time()
        ctx.upload_host_index(hidx)
        t_upload = time.time() - t0
    else:                                           # suffix sort, BWT, Occ blocks and SA on the device (index_dev.cu)
        hidx = ctx.build_index_device(contigs)
        t_index, t_upload = time.time() - t0, 0.0
    opt = abi.default_opt()

    # ---------------- device-resident timing ("value")
    ctx.batch_upload(seqs, off)
    for _ in range(a.warmup):
        ctx.batch_run(opt, want_stage_ms=True)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    stage = np.zeros(6)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        stage += np.array(ctx.batch_run(opt, want_stage_ms=True))
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.finish()
    cnt = ctx.counters()
    dev_ms = stage[5] / a.steps                       # CUDA-event bracket of the whole batch on the launching stream
    regs_total = cnt["regs"]

    # ---------------- end-to-end through the C ABI with pinned host buffers ("e2e")
    pin_seq = torch.empty(len(seqs), dtype=torch.uint8).pin_memory()
    pin_seq.numpy()[:] = seqs
    cap = int(regs_total * 1.1) + 1024
    pin_out = torch.empty(cap * abi.ALNREG_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
    pin_n = torch.empty(a.reads, dtype=torch.int32).pin_memory()
    pin_off = torch.empty(a.reads + 1, dtype=torch.int64).pin_memory()
    pin_off.numpy()[:] = off
    import ctypes as C

    def e2e_step():
        tot = ctx.lib.b200_align_batch(ctx.h, C.byref(opt), C.c_int(a.reads), C.c_void_p(pin_seq.data_ptr()),
                                       C.c_void_p(pin_off.data_ptr()), C.c_int64(cap),
                                       C.c_void_p(pin_out.data_ptr()), C.c_void_p(pin_n.data_ptr()))
        if tot < 0:
            raise SystemExit(f"b200_align_batch failed: {ctx.lib.b200_last_error().decode()} ({tot})")
        return tot

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        tot = e2e_step()
    barrier()
    e2e_wall = time.perf_counter() - t0

    # ---------------- max over ranks
    tvec = torch.tensor([dev_ms, wall * 1e3 / a.steps, e2e_wall * 1e3 / a.steps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tvec, op=dist.ReduceOp.MAX)
    dev_ms_max, wall_ms_max, e2e_ms_max = [float(x) for x in tvec.cpu()]

    if rank == 0:
        peak, peak_src = peaks()
        # dominant kernels: SMEM seeding = smem_p3_kernel (third pass) + smem_fast3_kernel (passes 1-2, merge, sort),
        # launched back to back once per chunk of 2^20 reads; bytes and time are per chunk, over both
        n_launch = max(1, -(-a.reads // (1 << 20)))
        alg_bytes = 64.0 * (cnt["extend"] + cnt["two_block"]) / n_launch      # counters are per batch_run
        smem_ms = stage[1] / a.steps
        achieved = alg_bytes / (smem_ms / n_launch * 1e-3) / 1e9 if smem_ms > 0 else 0.0
        mb = microbench_ceilings()
        ext_ms = stage[4] / a.steps
        out = {
            "metric": METRIC, "value": world * a.reads / (dev_ms_max * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": dev_ms_max, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": workload_name(a), "path": "mem_align1_core batch (reads -> mem_alnreg_t)",
                       "l2": "inputs + intermediates per step are GBs (>> 126 MB L2); no explicit flush",
                       "sharding": "reads sharded per rank, FM-index replicated, no collective on the data path",
                       "wall_ms_per_step": wall_ms_max, "index": "host SA-IS + upload" if a.index == "host" else "built on the device",
                       "index_build_s": t_index, "index_upload_s": t_upload, "seq_len": hidx.seq_len, "read_sim_s": t_reads},
            "e2e": {"value": world * a.reads / (e2e_ms_max * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": world * int(len(seqs) + off.nbytes),          # all ranks (each copies its own block)
                    "d2h_bytes_per_step": world * int(tot * abi.ALNREG_DTYPE.itemsize + 4 * a.reads)},
            "gpu_launches": int(cnt["launches"]) * a.steps,
            "clocks": clocks,
            "stage_ms": {k: float(v) / a.steps for k, v in zip(["pack", "smem", "seed_sa", "chain", "extend", "total"], stage)},
            "counters": cnt,
            "roofline": {"bound": "hbm", "kernel": SEED_KERNELS, "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak,
                         "traffic": measured_traffic(SEED_KERNELS) if (a.contigs, a.contig_len, a.read_len) == (4, 25_000_000, 150) else None,
                         "peak_source": peak_src, "algorithmic_bytes": alg_bytes, "launches_per_step": n_launch,
                         "ms_per_launch": smem_ms / n_launch,
                         "note": "algorithmic bytes = 64 B x Occ blocks touched by bwt_extend (1 or 2 per call, "
                                 "SURVEY.md 8d) per launch (one chunk of 2^20 reads; the seeding stage is the pair of kernels "
                                 "named in `kernel`, timed together); the kernels read one 32-B "
                                 "sector per Occ block from its own index layout and answers calls on strings of <= 9 bases from a "
                                 "k-mer interval table, so DRAM traffic is below that"},
            "extension": {"gcups": cnt["cells"] / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else None,
                          "cells": cnt["cells"], "ksw_calls": cnt["ksw"], "dpx_ops_per_cell": 5,
                          "dpx_peak_gcups": (mb["dpx_tops"] * 1e3 / 5) if mb and "dpx_tops" in mb else None,
                          "frac_of_dpx_peak": (cnt["cells"] / (ext_ms * 1e-3) / 1e9) / (mb["dpx_tops"] * 1e3 / 5)
                                              if mb and "dpx_tops" in mb and ext_ms > 0 else None},
            "gather_ceiling": ({"sectors_per_s_l2": mb.get("gather_l2_gsectors"), "sectors_per_s_dram": mb.get("gather_dram_gsectors"),
                                "achieved_gsectors": (cnt["extend"] + cnt["two_block"]) / (smem_ms * 1e-3) / 1e9 if smem_ms > 0 else None}
                               if mb else None),
            "smem_lookups_per_sec": cnt["extend"] / (smem_ms * 1e-3) if smem_ms > 0 else None,
        }
        if not a.no_cpu_baseline:
            n_sample = a.cpu_sample or min(a.reads, max(20000, 150000 * cores))
            chk = None
            if a.check > 0:
                n_chk = min(a.check, a.reads)
                dev_cnt = pin_n.numpy()[:n_chk].copy()
                dev_regs = np.frombuffer(pin_out.numpy()[: int(dev_cnt.sum()) * abi.ALNREG_DTYPE.itemsize].tobytes(), dtype=abi.ALNREG_DTYPE)
                chk = (n_chk, dev_regs, dev_cnt)
        all_cnt = pin_n.numpy()[: a.reads]
        all_regs = np.frombuffer(pin_out.numpy()[: int(all_cnt.astype(np.int64).sum()) * abi.ALNREG_DTYPE.itemsize], dtype=abi.ALNREG_DTYPE)
        out["reads_to_sam"], sam_chk = reads_to_sam(a, ctx, hidx, opt, pin_seq, pin_off, cores)
        if not a.no_cpu_baseline:
            cb = cpu_reference(a, hidx, seqs, off, n_sample, cores, chk, sam_chk)
            if cb is not None:
                out["parity_sample"] = cb.pop("parity_sample", None)
                out["reads_to_sam"]["parity_sample"] = cb.pop("sam_parity_sample", None)
                out["cpu_baseline"] = cb
        out["truth_check"] = truth_check(a, contigs, build_inputs.truth, all_regs, all_cnt)
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
