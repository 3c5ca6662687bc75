#!/usr/bin/env python
"""bench.py -- headline benchmark: aligned reads/s of the BWA-MEM alignment core on B200.

Workload (BASELINE.json configs[3], the config the metric is quoted on): synthetic 3.1 Gbp reference (24 contigs,
i.i.d. uniform bases, seed 20261104; 6.2 G-row FM-index, 64-bit positions), ONE FIXED SET of --reads (default 100 M) x 150 bp
single-end reads (1 % substitutions, 0.015 %/base 1-bp indels) SHARDED over the N GPUs: strong scaling.

  value : whole-job reads/s of the hot path -- worker1 / mem_align1_core of mem_process_seqs for every read (SMEM seeding ->
          SA location -> chaining/filtering -> banded extension -> dedup; reads in, mem_alnreg_t records out) -- with each
          GPU's shard already resident in its HBM; CUDA events per device, max over devices.
  e2e   : the same read set through the call a host program makes, b200_reads_to_sam: reads in pinned HOST memory -> SAM
          text in pinned HOST memory, merged in input order (worker1 + worker2 + SAM text on the devices, H2D and D2H
          inside the timed region).
  --impl reference : the reference's own CPU code (oracle/_ref, all host threads) on a bounded sample of the same read
          set and the same index: value = mem_align1_core over kt_for (worker1), e2e = the whole mem_process_seqs
          (bwamem.c:1215-1244: worker1 + worker2, reads -> SAM text).

One process drives all N GPUs (one host thread per device, csrc/reads2sam.cu) because the reference's CLI is one process;
under torchrun rank 0 does that and the other ranks only take part in the start / end barriers.  The index arrays are made
once per box by the device builder and cached as files (bwa_mem_gpu_b200/bench_index.py), so the reference arm never loads
the product library.
"""
from __future__ import annotations

import argparse
import ctypes as C
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
READ_SEED = 1004


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--reads", type=int, default=100_000_000, help="reads of the fixed set (whole job, all GPUs)")
    ap.add_argument("--contigs", type=int, default=24)
    ap.add_argument("--contig-len", type=int, default=129_166_667)
    ap.add_argument("--small", action="store_true", help="BASELINE.json configs[1]: 4 x 25 Mbp reference, 10 M reads")
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("--cpu-sample", type=int, default=0, help="reads in the CPU sample (0 = auto: about 4 s of CPU work per measurement)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="skip the paired-end and long-read sub-records")
    ap.add_argument("--tune", default="", help="comma-separated key=value context tunables (b200_ctx_tune)")
    ap.add_argument("--sub-reads", type=int, default=0, help="e2e: reads per sub-batch of b200_reads_to_sam (0 = the library's rule)")
    ap.add_argument("--ctx-per-gpu", type=int, default=3, help="e2e: host threads (contexts sharing one index) per device")
    ap.add_argument("--check", type=int, default=20000, help="reads compared with the reference CPU code (regions and SAM text)")
    a = ap.parse_args()
    if a.small:
        a.contigs, a.contig_len = 4, 25_000_000
        if a.reads == 100_000_000:
            a.reads = 10_000_000
    return a


def workload_name(a):
    return (f"synthetic {a.contigs}x{a.contig_len / 1e6:.6g} Mbp reference (seed 20261104), one fixed set of "
            f"{a.reads / 1e6:g}M x {a.read_len} bp single-end reads (1% sub + 0.015%/bp indel, seed {READ_SEED}) sharded over the GPUs")


def config_of(a):
    """The same keys and values in both arms."""
    return {"workload": workload_name(a),
            "path": "value: mem_align1_core batch (reads -> mem_alnreg_t), shards resident in HBM; e2e: reads in host memory -> SAM text "
                    "in host memory (mem_process_seqs: worker1 + worker2), input order",
            "l2": "inputs + intermediates per step are GBs (>> 126 MB L2); no explicit flush",
            "sharding": "fixed read set sharded over the GPUs, FM-index replicated, no collective on the data path",
            "reads_total": a.reads}


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


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


SEED_KERNELS = "smem_p3_kernel+smem_fast3_kernel"


def measured_traffic(kernel, key):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (profiles/traffic.json:
    dram__bytes_read.sum + dram__bytes_write.sum, one launch = one 2^20-read chunk) for this index size, or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    tab = json.load(open(p)).get(key)
    if not tab:
        return None
    tot = 0.0
    for k in kernel.split("+"):
        if k not in tab:
            return None
        tot += float(tab[k]["dram_bytes_per_launch"])
    return tot


def microbench_ceilings():
    p = os.path.join(ROOT, "profiles", "microbench.json")
    return json.load(open(p)) if os.path.exists(p) else None


def ensure_index_cache(a, device=0):
    """Cache directory with pac / bwt / sa of the bench genome; built in a SUBPROCESS when missing (the reference arm must
    not load the product library itself)."""
    from bwa_mem_gpu_b200 import bench_index as BI
    d = BI.cache_dir(ROOT, a.contigs, a.contig_len)
    if not BI.have(d):
        t0 = time.time()
        r = subprocess.run([sys.executable, "-m", "bwa_mem_gpu_b200.bench_index", "--contigs", str(a.contigs), "--contig-len",
                            str(a.contig_len), "--out", d, "--device", str(device)], cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        if r.returncode != 0 or not BI.have(d):
            raise SystemExit(f"index cache build failed: {r.stderr[-2000:]}")
        sys.stderr.write(f"[bench] index cache built in {time.time() - t0:.1f} s: {d}\n")
    return d


def truth_check(a, l_pac, truth, regs, cnt):
    """Size-independent property: a read drawn from a known place must come back there.  Fraction of reads whose best
    region (first record) lies on the true strand within 20 bp of the true start."""
    contig, pos, rev = truth
    n = len(cnt)
    first = np.concatenate([[0], np.cumsum(cnt.astype(np.int64))])[:-1]
    has = cnt > 0
    r = regs[first[has]]
    gpos = (contig * a.contig_len + pos)[:n][has]
    rv = (rev[:n] == 1)[has]
    on_rev = r["rb"] >= l_pac
    fwd_start = np.where(on_rev, 2 * l_pac - r["re"], r["rb"])
    clip = np.where(on_rev, a.read_len - r["qe"], r["qb"])
    ok = (on_rev == rv) & (np.abs(fwd_start - clip - gpos) <= 20)
    return {"reads": int(n), "with_region": int(has.sum()), "best_region_at_true_locus": int(ok.sum()), "frac": float(ok.sum()) / max(1, n)}


# ------------------------------------------------------------------------------------------------ reference arm
def ref_index_of(a):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refapi
    from bwa_mem_gpu_b200 import bench_index as BI
    if not refapi.available():
        return None, None, None
    d = ensure_index_cache(a)
    meta, pac, bwt, sa = BI.load(d, mmap=False)
    rl = refapi.RefLib()
    lens = np.full(a.contigs, a.contig_len, dtype=np.int64)
    ann_off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
    idx = rl.index_from_arrays(bwt, meta["primary"], meta["L2"], meta["seq_len"], sa, meta["sa_intv"], pac, meta["l_pac"], ann_off,
                               lens.astype(np.int32))
    return idx, meta, pac


def ref_sample(a, pac, meta, n_sample):
    from bwa_mem_gpu_b200 import bench_index as BI, synth
    g = BI.unpack_pac_numpy(pac, 0, meta["l_pac"])
    parts, CH = [], 200_000
    for b in range(0, n_sample, CH):
        parts.append(synth.counter_reads(np, g, a.contigs, a.contig_len, b, min(n_sample, b + CH), a.read_len, READ_SEED)[0])
    reads = np.concatenate(parts)
    return np.ascontiguousarray(reads.reshape(-1)), np.arange(0, n_sample + 1, dtype=np.int64) * a.read_len


def cpu_measure(idx, opt, seqs, off, threads):
    """One step of the reference's CPU code on the sample: worker1 alone (mem_align1_core over kt_for) and the whole
    mem_process_seqs (worker1 + worker2 + SAM text)."""
    from bwa_mem_gpu_b200 import synth
    n = len(off) - 1
    t0 = time.time()
    _, cnt = idx.align(opt, seqs, off, n_threads=threads, count_only=True)
    t1 = time.time()
    o2 = type(opt).from_buffer_copy(opt)
    o2.n_threads = threads
    sam = idx.process_seqs(o2, synth.to_ascii(seqs), off, n_processed=0)
    dt2 = idx.last_process_seqs_s                           # the mem_process_seqs call itself (its own "real sec"), not the marshalling around it
    return {"worker1_reads_per_s": n / (t1 - t0), "process_seqs_reads_per_s": n / dt2, "worker1_s": t1 - t0,
            "process_seqs_s": dt2, "regions": int(cnt.sum()), "sam_bytes": len(sam)}, sam


def auto_sample(a, cores):
    return a.cpu_sample or min(a.reads, max(20000, 40000 * cores))


def run_reference(a, cores):
    from bwa_mem_gpu_b200 import abi
    idx, meta, pac = ref_index_of(a)
    if idx is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libbwaref.so not built"}))
        return
    n_sample = auto_sample(a, cores)
    seqs, off = ref_sample(a, pac, meta, n_sample)
    opt = abi.default_opt()
    vals = []
    for s in range(a.warmup + a.steps):
        r, _ = cpu_measure(idx, opt, seqs, off, cores)
        if s >= a.warmup:
            vals.append(r)
    v1 = float(np.mean([x["worker1_reads_per_s"] for x in vals]))
    v2 = float(np.mean([x["process_seqs_reads_per_s"] for x in vals]))
    sample = (f"first {n_sample} reads of the fixed set per step; value: mem_align1_core via kt_for ({vals[-1]['worker1_s']:.2f} s), "
              f"e2e: mem_process_seqs, reads -> SAM text ({vals[-1]['process_seqs_s']:.2f} s); {cores} threads")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v1, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": 1e3 * n_sample / v1, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32",
        "data": "synthetic", "config": config_of(a),
        "cpu_baseline": {"value": v1, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample, "e2e_value": v2},
        "e2e": {"value": v2, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------------------------------------ our arm
def pinned(torch, nbytes):
    return torch.empty(int(nbytes), dtype=torch.uint8).pin_memory()


def make_reads(a, torch, BI, pac, l_pac, seq_pin):
    """The fixed read set, made on GPU 0 chunk by chunk into the pinned host buffer; truth of the first 2^20 reads."""
    dev = torch.device("cuda", 0)
    from bwa_mem_gpu_b200 import synth
    g = BI.unpack_pac_torch(torch.from_numpy(np.array(pac)).to(dev), l_pac)
    CH = 1 << 21
    view = seq_pin.view(-1, a.read_len)
    truth = None
    for b in range(0, a.reads, CH):
        e = min(a.reads, b + CH)
        r, contig, pos, rev = synth.counter_reads(torch, g, a.contigs, a.contig_len, b, e, a.read_len, READ_SEED)
        view[b:e].copy_(r, non_blocking=False)
        if b == 0:
            m = min(e, 1 << 20)
            truth = (contig[:m].cpu().numpy(), pos[:m].cpu().numpy(), rev[:m].cpu().numpy())
        del r, contig, pos, rev
    del g
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    return truth


def run_ours(a, world, cores, barrier):
    import torch
    from bwa_mem_gpu_b200 import abi, bench_index as BI, lib as L
    N = world
    if torch.cuda.device_count() < N:
        raise SystemExit(f"bench.py --gpus {N}: only {torch.cuda.device_count()} CUDA devices visible")
    t_setup = time.time()
    d = ensure_index_cache(a)
    if not a.no_sub and N == 1:                            # the paired-end sub-record's index too, while the device is still empty
        try:
            import bench_pe
            ensure_index_cache(argparse.Namespace(contigs=bench_pe.CONTIGS, contig_len=bench_pe.CONTIG_LEN))
        except (Exception, SystemExit) as e:  # noqa: BLE001
            sys.stderr.write(f"[bench] sub-record index: {str(e)[:300]}\n")
    meta, pac, bwt_words, sa_sampled = BI.load(d)
    l_pac = meta["l_pac"]
    # ---- inputs: reads and names in pinned host memory
    seq_pin = pinned(torch, a.reads * a.read_len)
    seqs = seq_pin.numpy()
    names_box = {}
    th_names = threading.Thread(target=lambda: names_box.update(v=L.numbered_names(b"r", 0, a.reads)))
    th_names.start()
    t0 = time.time()
    truth = make_reads(a, torch, BI, pac, l_pac, seq_pin)
    t_reads = time.time() - t0
    off_pin = pinned(torch, (a.reads + 1) * 8)
    off = off_pin.numpy().view(np.int64)
    off[:] = np.arange(0, a.reads + 1, dtype=np.int64) * a.read_len
    th_names.join()
    nf0, no0 = names_box["v"]
    nf_pin, no_pin = pinned(torch, len(nf0)), pinned(torch, no0.nbytes)
    nf, no = nf_pin.numpy(), no_pin.numpy().view(np.int64)
    nf[:] = nf0; no[:] = no0
    del nf0, no0
    # ---- one context per device: the index is built on device 0 from the cached .pac and fanned out device to device
    #      (b200_index_replicate: peer copies over NVLink); a device whose peer copy fails builds its own
    ctxs, errs = [None] * N, []

    def tuned(dv):
        c = L.Context(dv)
        for kv in filter(None, a.tune.split(",")):
            k, v = kv.split("=")
            c.tune(k, int(v))
        return c
    t0 = time.time()
    ctxs[0] = tuned(0)
    hidx = ctxs[0].build_index_device_pac(np.asarray(pac), [a.contig_len] * a.contigs)
    t_index = time.time() - t0

    def mk(dv):
        try:
            c = tuned(dv)
            try:
                c.replicate_index_from(ctxs[0])
            except L.B200Error:
                c.build_index_device_pac(np.asarray(pac), [a.contig_len] * a.contigs)
            ctxs[dv] = c
        except Exception as e:  # noqa: BLE001
            errs.append(f"device {dv}: {e}")
    t0 = time.time()
    ths = [threading.Thread(target=mk, args=(dv,)) for dv in range(1, N)]
    [t.start() for t in ths]; [t.join() for t in ths]
    if errs:
        raise SystemExit("; ".join(errs))
    t_fanout = time.time() - t0
    opt = abi.default_opt()
    bounds = [(int(a.reads * dv // N), int(a.reads * (dv + 1) // N)) for dv in range(N)]

    # ---------------- device-resident timing ("value"): every device its shard, K steps back to back, CUDA events
    stage = np.zeros((N, 6))
    cnts = [None] * N
    bar = threading.Barrier(N)

    def resident(dv):
        try:
            lo, hi = bounds[dv]
            c = ctxs[dv]
            c.batch_upload(seqs, off[lo: hi + 1])
            # The algorithmic traffic of seeding (bwt_extend calls, two-block calls: a property of the input) is counted
            # by the instrumented template of the seeding kernels in one untimed pass; every other launch, the timed
            # ones included, runs the kernels without the counter arithmetic (tunable seed_stats).
            c.tune("seed_stats", 1)
            c.batch_run(opt, want_stage_ms=True)
            cw = c.counters()
            c.tune("seed_stats", 0)
            for _ in range(a.warmup):
                c.batch_run(opt, want_stage_ms=True)
            bar.wait()
            for _ in range(a.steps):
                stage[dv] += np.array(c.batch_run(opt, want_stage_ms=True))
            cn = c.counters()
            cn["extend"], cn["two_block"] = cw["extend"], cw["two_block"]
            cnts[dv] = cn
        except Exception as e:  # noqa: BLE001
            errs.append(f"device {dv}: {e}")
            bar.abort()
    barrier()
    sampler = ClockSampler(0)
    sampler.start()
    t0 = time.perf_counter()
    ths = [threading.Thread(target=resident, args=(dv,)) for dv in range(N)]
    [t.start() for t in ths]; [t.join() for t in ths]
    if errs:
        raise SystemExit("; ".join(errs))
    wall_value = time.perf_counter() - t0
    dev_ms = float(stage[:, 5].max()) / a.steps        # max over devices of the CUDA-event bracket of the whole shard
    # parity / truth inputs while device 0 still holds the results of its shard
    cnt0 = cnts[0]

    # ---------------- end to end ("e2e"): pinned reads -> pinned SAM text through b200_reads_to_sam over all devices
    sam_cap = int(a.reads * (2 * a.read_len + 110)) + (1 << 20)
    sam_pin = pinned(torch, sam_cap)
    sam = sam_pin.numpy()
    e2e_s, e2e_stats, sam_bytes = [], None, 0
    e2e_ctxs = []                                          # --ctx-per-gpu host threads per device, each with its own context borrowing the device's index
    for dv in range(N):
        e2e_ctxs.append(ctxs[dv])
        for _ in range(max(1, a.ctx_per_gpu) - 1):
            c2 = L.Context(dv)
            for kv in filter(None, a.tune.split(",")):
                k, v = kv.split("=")
                c2.tune(k, int(v))
            c2.replicate_index_from(ctxs[dv])
            e2e_ctxs.append(c2)
    for s in range(a.warmup + a.steps):
        t0 = time.perf_counter()
        out, _, st = L.reads_to_sam(e2e_ctxs, opt, hidx.names, seqs, off, nf, no, sam=sam, want_off=False, sub_reads=a.sub_reads)
        dt = time.perf_counter() - t0
        if s >= a.warmup:
            e2e_s.append(dt)
            e2e_stats, sam_bytes = st, len(out)
    clocks = sampler.finish()
    e2e_ms = 1e3 * float(np.mean(e2e_s))
    # offsets of the records once, outside the timed region (parity sample below)
    out, _, _ = L.reads_to_sam(ctxs[:1], opt, hidx.names, seqs, off[: min(a.reads, max(a.check, 1)) + 1], nf, no, want_off=False)
    sam_head = out.tobytes()

    cnt = {k: sum(c[k] for c in cnts) for k in cnts[0]}
    peak, peak_src = peaks()
    mb = microbench_ceilings()
    chunk = 1 << 20
    n_launch = sum(max(1, -(-(hi - lo) // chunk)) for lo, hi in bounds)           # seeding launches (kernel pairs) per step, all devices
    smem_ms = float(stage[:, 1].sum()) / a.steps                                  # summed over devices
    alg_bytes = 64.0 * (cnt["extend"] + cnt["two_block"]) / n_launch
    ms_per_launch = smem_ms / n_launch
    achieved = alg_bytes / (ms_per_launch * 1e-3) / 1e9 if smem_ms > 0 else 0.0
    ext_ms = float(stage[:, 4].sum()) / a.steps
    gcups = cnt["cells"] / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else None
    dpx_peak = (mb["dpx_tops"] * 1e3 / 5) if mb and "dpx_tops" in mb else None
    key = "3.1Gbp" if a.contigs * a.contig_len > 2_000_000_000 else "100Mbp"
    roof_seed = {"bound": "hbm", "kernel": SEED_KERNELS, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                 "traffic": measured_traffic(SEED_KERNELS, key), "peak_source": peak_src, "algorithmic_bytes": alg_bytes,
                 "launches_per_step": n_launch, "ms_per_launch": ms_per_launch,
                 "note": "algorithmic bytes = 64 B x Occ blocks touched by bwt_extend (1 or 2 per call, SURVEY.md 8d) per launch (one chunk of "
                         "up to 2^20 reads; the stage is the pair of kernels named in `kernel`, timed together, summed over the devices); "
                         "the kernels read one 32-B sector per Occ block from their own index layout and answer calls on short strings from "
                         "a k-mer interval table, so DRAM traffic is below that"}
    out_rec = {
        "metric": METRIC, "value": a.reads / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": N, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": config_of(a),
        "e2e": {"value": a.reads / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(len(seqs) + off.nbytes + len(nf) + no.nbytes),
                "d2h_bytes_per_step": int(sam_bytes), "ms_per_step": e2e_ms, "sam_bytes": int(sam_bytes),
                "call": f"b200_reads_to_sam (one process, {max(1, a.ctx_per_gpu)} host threads per device sharing its index; worker1 + worker2 + SAM text on the devices; text placed in input order)",
                "thread_seconds": {"worker1": e2e_stats[1], "worker2_and_text": e2e_stats[2], "placement_wait": e2e_stats[3]},
                "sub_batches": int(e2e_stats[4])},
        "gpu_launches": int(cnt["launches"]) * a.steps,
        "clocks": clocks,
        "stage_ms": {k: float(v) / a.steps for k, v in zip(["pack", "smem", "seed_sa", "chain", "extend", "total"], stage.max(axis=0))},
        "counters": cnt,
        "roofline": roof_seed,
        "rooflines": [roof_seed,
                      {"bound": "dpx", "kernel": "ext_dp8_kernel (ksw_extend2)", "achieved": gcups, "peak": dpx_peak, "unit": "GCUPS",
                       "frac": (gcups / dpx_peak) if gcups and dpx_peak else None, "cells_per_step": cnt["cells"], "ksw_calls": cnt["ksw"],
                       "dpx_ops_per_cell": 5, "peak_source": "profiles/microbench.json: measured DPX issue rate / 5 instructions per cell",
                       "ms_per_step_all_devices": ext_ms}],
        "smem_lookups_per_sec": cnt["extend"] / (smem_ms * 1e-3) * N if smem_ms > 0 else None,
        "setup": {"index": "cached .pac -> b200_index_build_device on device 0, b200_index_replicate (peer copy) to the others", "index_build_s": t_index,
                  "index_fanout_s": t_fanout, "read_sim_s": t_reads,
                  "seq_len": hidx.seq_len, "wall_value_phase_s": wall_value, "total_wall_s": None},
    }
    # ---------------- checks against the reference CPU code on a sample (and the CPU baseline at N = 1)
    n_chk = min(a.check, a.reads)
    if n_chk > 0:
        regs_h, cnt_h = ctxs[0].align(opt, seqs, off[: min(a.reads, 1 << 20) + 1])
        out_rec["truth_check"] = truth_check(a, l_pac, truth, regs_h, cnt_h)
        if not a.no_cpu_baseline:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import refapi
            if refapi.available():
                rl = refapi.RefLib()
                lens = np.full(a.contigs, a.contig_len, dtype=np.int64)
                ann_off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
                idx = rl.index_from_arrays(np.asarray(bwt_words), meta["primary"], meta["L2"], meta["seq_len"], np.asarray(sa_sampled),
                                           meta["sa_intv"], np.asarray(pac), l_pac, ann_off, lens.astype(np.int32))
                r0, c0 = idx.align(opt, seqs, off[: n_chk + 1], n_threads=cores, cap_per_read=64)
                k = int(c0.sum())
                same = bool(np.array_equal(c0, cnt_h[:n_chk])) and r0.tobytes() == regs_h[:k].tobytes()
                o2 = abi.default_opt(); o2.n_threads = cores
                from bwa_mem_gpu_b200 import synth
                sam0 = idx.process_seqs(o2, synth.to_ascii(seqs[: n_chk * a.read_len]), off[: n_chk + 1], n_processed=0)
                out_rec["parity_sample"] = {"reads": int(n_chk), "regions": k, "regions_identical_to_reference": same,
                                            "sam_bytes": len(sam0), "sam_identical_to_reference": bool(sam0 == sam_head),
                                            "sam_of_timed_run_identical": bool(sam0 == sam[: len(sam0)].tobytes())}
                if N == 1:
                    n_sample = auto_sample(a, cores)
                    r, _ = cpu_measure(idx, opt, seqs[: n_sample * a.read_len], off[: n_sample + 1], cores)
                    out_rec["cpu_baseline"] = {"value": r["worker1_reads_per_s"], "unit": UNIT, "cores": cores, "kind": "reference",
                                               "sample": f"first {n_sample} reads of the fixed set: mem_align1_core via kt_for {r['worker1_s']:.2f} s; "
                                                         f"mem_process_seqs (reads -> SAM) {r['process_seqs_s']:.2f} s",
                                               "e2e_value": r["process_seqs_reads_per_s"]}
                idx = None
    if not a.no_sub and N == 1:
        for c in e2e_ctxs[::-1]:                           # the sub-records need the device memory: keep only the index owner
            if c is not ctxs[0]:
                c.close()
        ctxs[0].trim()                                     # ... without the 100 M-read batch it still holds
        try:
            import bench_sub
            out_rec["sub_records"] = bench_sub.run(a, ctxs[0], hidx, pac, cores)
        except (Exception, SystemExit) as e:  # noqa: BLE001  (a sub-record must never cost the main line)
            out_rec["sub_records"] = {"error": str(e)[:500]}
    out_rec["setup"]["host_cpus"] = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    out_rec["setup"]["total_wall_s"] = time.time() - t_setup
    print(json.dumps(out_rec))
    for c in e2e_ctxs[::-1]:
        c.close()


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = os.cpu_count() or 1
    if a.impl == "reference":
        if rank == 0:
            run_reference(a, cores)
        return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU implementation")
    dist = None
    cpu_group = None
    if world > 1:
        import datetime
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        # keep stdout to the one JSON line: NCCL writes its version banner to fd 1 when the communicator is created
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank), timeout=datetime.timedelta(hours=2))
            dist.barrier()                                 # one NCCL round over all ranks; everything after waits on the CPU (gloo):
            torch.cuda.synchronize()                       # a rank parked in an NCCL kernel would spin on the GPU rank 0 is driving
            cpu_group = dist.new_group(backend="gloo", timeout=datetime.timedelta(hours=2))
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier(group=cpu_group)
    n_dev = world if world > 1 else a.gpus                 # plain `python bench.py --gpus N` drives N devices too
    if rank == 0:
        run_ours(a, n_dev, cores, barrier)
    else:
        barrier()                                          # the start barrier of the timed region; rank 0 drives every device from here
    if dist is not None:
        dist.barrier(group=cpu_group)                      # end of job: everybody leaves together
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
