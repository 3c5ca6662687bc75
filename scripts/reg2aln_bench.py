"""mem_reg2aln on the device (b200_reg2aln_batch): regions of N reads of the configs[1] workload ->
positions, CIGARs, NM, MD.  Times the whole host-buffer call (uploads, kernels, compaction, downloads) and the reference's
mem_reg2aln on ONE host thread over a sample (the shim loop is serial); checks the sample for identity.
usage: python scripts/reg2aln_bench.py [--reads N] [--check M]"""
import argparse, json, os, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from bwa_mem_gpu_b200 import abi, lib as L, synth
import refapi


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=2_000_000)
    ap.add_argument("--check", type=int, default=100_000)
    a = ap.parse_args()
    contigs = synth.make_genome([25_000_000] * 4)
    reads, _ = synth.simulate_reads(contigs, a.reads, seed=1002)
    seqs = np.ascontiguousarray(reads.reshape(-1)); off = np.arange(0, a.reads + 1, dtype=np.int64) * 150
    ctx = L.Context(0)
    hidx = ctx.build_index_device(contigs)
    opt = abi.default_opt()
    regs, cnt = ctx.align(opt, seqs, off)
    ctx.reg2aln(opt, seqs[: 150 * 1000], off[:1001], regs[: int(cnt[:1000].sum())], cnt[:1000])      # warm-up
    ctx.reg2aln(opt, seqs, off, regs, cnt)                                                            # sizes the buffers
    t0 = time.time()
    aln, cig, md = ctx.reg2aln(opt, seqs, off, regs, cnt)
    dt = time.time() - t0
    res = {"reads": a.reads, "regions": int(len(regs)), "device_call_s": dt, "regions_per_s": len(regs) / dt,
           "cigar_words": int(len(cig)), "md_bytes": int(len(md)), "NM_mean": float(aln["NM"].mean())}
    if refapi.available():
        m = min(a.check, a.reads)
        rl = refapi.RefLib()
        ridx = rl.index_from_arrays(hidx.bwt_words(), hidx.primary, hidx.L2, hidx.seq_len, hidx.sa(), hidx.sa_intv,
                                    hidx.pac(), hidx.l_pac, hidx.ann_off, hidx.ann_len)
        k = int(cnt[:m].sum())
        t0 = time.time()
        a0, c0, m0 = ridx.reg2aln(opt, seqs, off[: m + 1], regs[:k], cnt[:m])
        t_ref = time.time() - t0
        nc = int(aln["n_cigar"][:k].sum()); nm = int(aln["l_md"][:k].sum())
        same = (np.array_equal(a0[:, 0], aln["pos"][:k]) and np.array_equal(a0[:, 3], aln["n_cigar"][:k]) and
                np.array_equal(a0[:, 4], aln["NM"][:k]) and np.array_equal(c0, cig[:nc]) and m0.tobytes() == md[:nm].tobytes())
        res.update({"ref_sample_regions": k, "ref_regions_per_s_1thread": k / t_ref, "identical_to_reference": bool(same),
                    "host_cores": os.cpu_count()})
    print(json.dumps(res))


if __name__ == "__main__":
    main()
