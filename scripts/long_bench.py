"""Long-read measurement (BASELINE.json configs[4] shape on the 100 Mbp reference): reads of 1-10 kbp with 3 %
substitutions and 4 % indels; device pipeline time, per-stage times and the ksw_align2 / ksw_extend2 work counters;
parity of a sample against the reference's mem_align1_core.  Measurement helper, not the product path.
usage: python scripts/long_bench.py [--reads N] [--check M]"""
import argparse, json, os, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from bwa_mem_gpu_b200 import abi, lib as L, synth
import refapi


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=20000)
    ap.add_argument("--check", type=int, default=300)
    ap.add_argument("--contigs", type=int, default=4)
    ap.add_argument("--contig-len", type=int, default=25_000_000)
    ap.add_argument("--chunk", type=int, default=0, help="reads per pipeline chunk (0 = the library's own choice)")
    ap.add_argument("--tune", default="", help="comma-separated key=value context tunables")
    ap.add_argument("--sam", action="store_true", help="also worker2 on the resident batch + SAM text (single-end), with a sample check")
    a = ap.parse_args()
    contigs = synth.make_genome([a.contig_len] * a.contigs)
    rng = np.random.default_rng(5)
    lens = rng.integers(1000, 10001, size=a.reads)
    seqs, off = [], [0]
    for ln in sorted(set(int(x) // 500 * 500 + 500 for x in lens)):          # simulate per length class, then trim
        idx = np.nonzero((lens // 500 * 500 + 500) == ln)[0]
        r, _ = synth.simulate_reads(contigs, len(idx), read_len=ln, seed=1000 + ln, sub_rate=0.03, indel_rate=0.04)
        for k, i in enumerate(idx):
            seqs.append((int(i), r[k, : int(lens[i])]))
    seqs.sort(key=lambda t: t[0])
    flat = np.ascontiguousarray(np.concatenate([s for _, s in seqs]).astype(np.uint8))
    off = np.concatenate([[0], np.cumsum([len(s) for _, s in seqs])]).astype(np.int64)
    ctx = L.Context(0)
    if a.chunk > 0:
        ctx.tune("chunk_reads", a.chunk)
    for kv in filter(None, a.tune.split(",")):
        k, v = kv.split("=")
        ctx.tune(k, int(v))
    hidx = ctx.build_index_device(contigs)
    opt = abi.default_opt()
    ctx.batch_upload(flat, off)
    ctx.batch_run(opt)
    t0 = time.time()
    ms = ctx.batch_run(opt, want_stage_ms=True)
    wall = time.time() - t0
    cnt = ctx.counters()
    res = {"chunk_reads": a.chunk, "tune": a.tune, "reads": a.reads, "bases": int(off[-1]), "device_ms": ms[5], "wall_s": wall,
           "stage_ms": dict(zip(["pack", "smem", "seed_sa", "chain", "extend", "total"], ms)),
           "reads_per_s": a.reads / (ms[5] * 1e-3), "mbases_per_s": off[-1] / (ms[5] * 1e-3) / 1e6,
           "ksw_extend2_gcups": cnt["cells"] / (ms[4] * 1e-3) / 1e9, "counters": cnt}
    m = min(a.check, a.reads)
    regs, n = ctx.batch_download(a.reads, 16 * a.reads)
    if refapi.available():
        rl = refapi.RefLib()
        ridx = rl.index_from_arrays(hidx.bwt_words(), hidx.primary, hidx.L2, hidx.seq_len, hidx.sa(), hidx.sa_intv,
                                    hidx.pac(), hidx.l_pac, hidx.ann_off, hidx.ann_len)
        t0 = time.time()
        r0, n0 = ridx.align(opt, flat, off[: m + 1], n_threads=os.cpu_count() or 1, cap_per_read=512)
        t_ref = time.time() - t0
        k = int(n[:m].sum())
        res.update({"parity_sample": m, "parity": bool(np.array_equal(n0, n[:m]) and r0.tobytes() == regs[:k].tobytes()),
                    "ref_reads_per_s_all_cores": m / t_ref, "ref_cores": os.cpu_count()})
    if a.sam:                                              # worker2 on the resident batch (b200_batch_reg2sam_se) + b200_sam_se_format
        os.environ["B200_R2A_TIMING"] = "1"
        ctx.batch_reg2sam_se(opt, n)                       # sizes the buffers
        t0 = time.time()
        r2, sel, aln, cig, md = ctx.batch_reg2sam_se(opt, n)
        t_w2 = time.time() - t0
        names = [f"r{i}" for i in range(a.reads)]
        t0 = time.time()
        sam, soff = ctx.sam_se(opt, hidx.names, names, flat, off, n, sel, aln, cig, md, n_threads=os.cpu_count() or 8)
        t_fmt = time.time() - t0
        res.update({"worker2_call_s": t_w2, "sam_format_s_incl_python": t_fmt, "sam_bytes": len(sam), "regions": int(len(regs)),
                    "cigar_words": int(len(cig))})
        if refapi.available():
            lut = np.full(256, 4, dtype=np.uint8); lut[:4] = np.arange(4)
            k = int(n[:m].sum())
            t0 = time.time()
            _, _, sam0 = ridx.reg2sam_se(opt, lut[flat[: int(off[m])]], off[: m + 1], regs[:k], n[:m])
            t_ref = time.time() - t0
            res.update({"sam_parity": bool(sam0 == sam[: int(soff[m])]), "ref_worker2_reads_per_s_1thread": m / t_ref})
    print(json.dumps(res))


if __name__ == "__main__":
    main()
