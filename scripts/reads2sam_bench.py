"""Reads in pinned host memory -> SAM text in pinned host memory through b200_reads_to_sam (worker1 + worker2 + SAM text on
the device(s), csrc/reads2sam.cu), next to the host-formatter call b200_align_sam_se_batch on the same reads; the two
texts must be identical.  usage: python scripts/reads2sam_bench.py [--human] [--reads N] [--gpus G] [--host-threads T]"""
import argparse, ctypes as C, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bwa_mem_gpu_b200 import abi, lib as L, synth
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--human", action="store_true")
ap.add_argument("--reads", type=int, default=4_000_000)
ap.add_argument("--gpus", type=int, default=1)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--host-threads", type=int, default=os.cpu_count() or 8)
ap.add_argument("--no-host", action="store_true")
ap.add_argument("--tune", default="")
ap.add_argument("--ctx-per-gpu", type=int, default=1, help="host threads (contexts sharing one index) per device")
a = ap.parse_args()
nc, cl = (24, 129_166_667) if a.human else (4, 25_000_000)
contigs = synth.make_genome([cl] * nc)
reads, _ = synth.simulate_reads(contigs, a.reads, seed=1002)
n = a.reads
pin = lambda nbytes: torch.empty(nbytes, dtype=torch.uint8).pin_memory().numpy()
seqs = pin(n * 150); seqs[:] = reads.reshape(-1); del reads
off = pin((n + 1) * 8).view(np.int64); off[:] = np.arange(0, n + 1, dtype=np.int64) * 150
nf0, no0 = L.numbered_names(b"r", 0, n)
nf = pin(len(nf0)); nf[:] = nf0; no = pin(no0.nbytes).view(np.int64); no[:] = no0
ctxs = []
for d in range(a.gpus):
    c = L.Context(d)
    for kv in filter(None, a.tune.split(",")):
        k, v = kv.split("="); c.tune(k, int(v))
    t0 = time.time(); hidx = c.build_index_device(contigs); t_idx = time.time() - t0
    ctxs.append(c)
    for _ in range(a.ctx_per_gpu - 1):
        c2 = L.Context(d)
        c2.replicate_index_from(c)
        ctxs.append(c2)
opt = abi.default_opt()
sam = pin(n * 420 + 4096)
best, st = None, None
for rep in range(a.reps + 1):
    t0 = time.perf_counter()
    out, soff, stats = L.reads_to_sam(ctxs, opt, hidx.names, seqs, off, nf, no, sam=sam, want_off=False)
    dt = time.perf_counter() - t0
    if rep and (best is None or dt < best):
        best, st = dt, stats
res = {"reads": n, "gpus": a.gpus, "ctx_per_gpu": a.ctx_per_gpu, "genome_mbp": nc * cl / 1e6, "index_s": t_idx, "device_formatter": {"seconds": best, "Mreads_s": n / best / 1e6, "sam_bytes": len(out),
       "thread_s": {"worker1": st[1], "worker2_format": st[2], "placement_wait": st[3]}, "sub_batches": st[4]}}
if not a.no_host:
    names = [b"r%d" % i for i in range(n)]
    r_names = (C.c_char_p * n)(*names)
    a_names = (C.c_char_p * len(hidx.names))(*[x.encode() for x in hidx.names])
    sam2 = pin(len(out) + 4096); soff2 = np.zeros(n + 1, dtype=np.int64)
    P = lambda x, ty: x.ctypes.data_as(C.POINTER(ty))
    bh = None
    for rep in range(3):
        t0 = time.perf_counter()
        nb = ctxs[0].lib.b200_align_sam_se_batch(ctxs[0].h, C.byref(opt), a_names, None, C.c_int(n), C.c_int64(0), r_names, None,
                                                 P(seqs, C.c_uint8), None, P(off, C.c_int64), None, C.c_int(a.host_threads), C.c_int(0),
                                                 C.c_int64(len(sam2)), C.c_void_p(sam2.ctypes.data), P(soff2, C.c_int64))
        dt = time.perf_counter() - t0
        if rep and (bh is None or dt < bh):
            bh = dt
    res["host_formatter_1gpu"] = {"seconds": bh, "Mreads_s": n / bh / 1e6, "threads": a.host_threads,
                                  "identical_text": bool(nb == len(out) and np.array_equal(sam2[:nb], out))}
print(json.dumps(res))
