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
    import ctypes as C
    P = lambda a, ty: a.ctypes.data_as(C.POINTER(ty))
    t = {}
    for rep in range(2):                                   # the first pass sizes the grow-only buffers
        t0 = time.time()
        regs, cnt = ctx.align(opt, seqs, off)
        t["align_s"] = time.time() - t0
    # worker2 on the device, timed around the C call alone (host buffers in, host buffers out)
    nr = len(regs)
    sel = np.zeros(nr + 1, dtype=abi.SEL_DTYPE); aln = np.zeros(nr + 1, dtype=abi.ALN_DTYPE)
    cig = np.zeros(4 * nr + 64, dtype=np.uint32); md = np.zeros(16 * nr + 64, dtype=np.uint8)
    tot2 = (C.c_int64 * 2)()
    for rep in range(2):
        regs2 = regs.copy()
        t0 = time.time()
        rc = ctx.lib.b200_reg2sam_se_batch(ctx.h, C.byref(opt), C.c_int(a.reads), C.c_int64(0), P(seqs, C.c_uint8), P(off, C.c_int64),
                                           C.c_void_p(regs2.ctypes.data), P(cnt, C.c_int32), C.c_void_p(sel.ctypes.data),
                                           C.c_void_p(aln.ctypes.data), C.c_int64(len(cig)), P(cig, C.c_uint32), C.c_int64(len(md)),
                                           C.c_void_p(md.ctypes.data), tot2)
        t["reg2sam_se_s"] = time.time() - t0
        assert rc == nr, (rc, ctx.lib.b200_last_error())
    a_names = (C.c_char_p * len(hidx.names))(*[x.encode() for x in hidx.names])
    r_names = (C.c_char_p * a.reads)(*[x.encode() for x in names])
    soff = np.zeros(a.reads + 1, dtype=np.int64)
    sam_buf = np.zeros(260 * a.reads + 4096, dtype=np.uint8)
    for rep in range(2):
        t0 = time.time()
        rc = ctx.lib.b200_sam_se_format(C.byref(opt), a_names, None, C.c_int(a.reads), r_names, None, P(seqs, C.c_uint8), None,
                                        P(off, C.c_int64), P(cnt, C.c_int32), C.c_void_p(sel.ctypes.data), C.c_void_p(aln.ctypes.data),
                                        P(cig, C.c_uint32), C.c_void_p(md.ctypes.data), None, C.c_int(a.threads),
                                        C.c_int64(len(sam_buf)), C.c_void_p(sam_buf.ctypes.data), P(soff, C.c_int64))
        t["sam_format_s"] = time.time() - t0
        assert rc > 0, rc
    sam = sam_buf[:rc].tobytes()
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
