"""Parity against the COMMITTED golden vectors (tests/golden/*.npz, produced from the unmodified reference CPU
code by tests/golden/make_golden.py).  Needs neither /root/reference nor oracle/_ref: inputs are re-derived
from seeds, the device index comes from our own host index builder.  Bit-exact comparisons only.
"""
import os

import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import abi

BACKENDS = [pytest.param("emu", id="emu"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]
G = "tiny"


def _load(name):
    return np.load(os.path.join(cm.GOLDEN, name))


@pytest.mark.parametrize("backend", BACKENDS)
def test_golden_fmindex(backend):
    g, ctx = _load("fmindex.npz"), cm.context(backend, G)
    h = ctx._host_index
    assert [h.primary, h.seq_len, h.bwt_size, h.n_sa] + h.L2 == [int(x) for x in g["meta"]]
    assert np.array_equal(ctx.occ4(g["k"]), g["occ4"])
    a, b = ctx.occ2x4(g["k"], g["l"])
    assert np.array_equal(a, g["occk"]) and np.array_equal(b, g["occl"])
    assert np.array_equal(ctx.sa_lookup(g["sa_k"], use_full=False), g["sa"])
    assert np.array_equal(ctx.sa_lookup(g["sa_k"], use_full=True), g["sa"])
    assert np.array_equal(ctx.extend(g["ik"], 0), g["ext_fwd"])
    assert np.array_equal(ctx.extend(g["ik"], 1), g["ext_back"])
    assert np.array_equal(ctx.extend(g["ik2"], 1), g["ext2_back"])


@pytest.mark.parametrize("backend", BACKENDS)
def test_golden_smem1a(backend):
    g, ctx = _load("smem1a.npz"), cm.context(backend, G)
    seqs, off = cm.reads_for(G, 300, seed=11, n_rate=0.004)
    for max_intv in (0, 20):
        out, cnt, ret = ctx.smem1a(seqs, off, g["xs"], g["mi"], max_intv=max_intv, cap=160)
        assert np.array_equal(cnt, g[f"cnt{max_intv}"]) and np.array_equal(ret, g[f"ret{max_intv}"])
        got = np.concatenate([out[r, :cnt[r]] for r in range(300)]).view(np.uint64).reshape(-1, 4)
        assert np.array_equal(got, g[f"iv{max_intv}"])


@pytest.mark.parametrize("backend", BACKENDS)
def test_golden_collect_intv(backend):
    g, ctx = _load("collect_intv.npz"), cm.context(backend, G)
    seqs, off = cm.reads_for(G, 600, seed=21, n_rate=0.001)
    iv, cnt = ctx.collect_intv(abi.default_opt(), seqs, off)
    assert np.array_equal(cnt, g["cnt"])
    assert np.array_equal(iv.view(np.uint64).reshape(-1, 4), g["iv"])


@pytest.mark.parametrize("backend", BACKENDS)
def test_golden_chains(backend):
    g, ctx = _load("chains.npz"), cm.context(backend, G)
    seqs, off = cm.reads_for(G, 500, seed=31)
    for stage in (0, 1):
        chn, nch, sd = ctx.chains(abi.default_opt(), seqs, off, stage=stage, cap_chn_per_read=256, cap_seed_per_read=4096)
        assert np.array_equal(nch, g[f"nch{stage}"])
        assert np.array_equal(chn, g[f"chn{stage}"]) and np.array_equal(sd, g[f"sd{stage}"])


@pytest.mark.parametrize("backend", BACKENDS)
def test_golden_ksw_extend2(backend):
    g, ctx = _load("ksw_extend2.npz"), cm.context(backend, G)
    opt = abi.default_opt()
    args = [g[k] for k in ("ql", "qoff", "q", "tl", "toff", "t")]
    assert np.array_equal(ctx.ksw_extend2(*args, opt, g["w"], 5, 100, g["h0"]), g["out_z100_eb5"])
    assert np.array_equal(ctx.ksw_extend2(*args, opt, g["w"], 0, 0, g["h0"]), g["out_z0_eb0"])


@pytest.mark.parametrize("backend", BACKENDS)
def test_golden_ksw_align2(backend):
    g, ctx = _load("ksw_align2.npz"), cm.context(backend, G)
    args = [g[k] for k in ("ql", "qoff", "q", "tl", "toff", "t")]
    assert np.array_equal(ctx.ksw_align2(*args, abi.default_opt(), g["xtra"]), g["out"])


@pytest.mark.parametrize("backend", BACKENDS)
def test_golden_align(backend):
    ctx = cm.context(backend, G)
    opt = abi.default_opt()
    g = _load("align.npz")
    seqs, off = cm.reads_for(G, 800, seed=41, indel=0.004)
    regs, n = ctx.align(opt, seqs, off)
    assert np.array_equal(n, g["n"])
    assert np.array_equal(regs.view(np.uint8).reshape(-1, abi.ALNREG_DTYPE.itemsize), g["regs"])
    g = _load("align_ragged.npz")
    seqs, off = cm.ragged_reads(G)
    regs, n = ctx.align(opt, seqs, off)
    assert np.array_equal(n, g["n"])
    assert np.array_equal(regs.view(np.uint8).reshape(-1, abi.ALNREG_DTYPE.itemsize), g["regs"])


@pytest.mark.parametrize("backend", BACKENDS)
def test_ctx_trim_keeps_the_index(backend):
    """b200_ctx_trim gives the batch pipeline's device buffers back; the index stays and the next call regrows them."""
    from bwa_mem_gpu_b200 import lib as L
    ctx = L.Context(0, cm.backend_lib(backend))
    ctx.upload_host_index(cm.context(backend, G)._host_index)
    opt = abi.default_opt()
    g = _load("align.npz")
    seqs, off = cm.reads_for(G, 800, seed=41, indel=0.004)
    for _ in range(2):
        regs, n = ctx.align(opt, seqs, off)
        assert np.array_equal(n, g["n"])
        assert np.array_equal(regs.view(np.uint8).reshape(-1, abi.ALNREG_DTYPE.itemsize), g["regs"])
        ctx.trim()
    ctx.close()


@pytest.mark.parametrize("backend", BACKENDS)
def test_full_sa_five_bytes_per_row(backend):
    """The full suffix array packed to 5 bytes per row (what an index of 2^32 or more rows gets; forced here with
    force_wide): every lookup equals the bwt_sa walk (bwt.c:86-96) and the golden values, including row 0, the last
    row, and rows at all four byte alignments of the packed words."""
    from bwa_mem_gpu_b200 import lib as L
    g = _load("fmindex.npz")
    ctx = L.Context(0, cm.backend_lib(backend))
    ctx.tune("force_wide", 1)
    h = cm.context(backend, G)._host_index
    ctx.upload_host_index(h)
    assert np.array_equal(ctx.sa_lookup(g["sa_k"], use_full=True), g["sa"])
    rng = np.random.default_rng(5)
    k = np.concatenate([np.arange(0, 64, dtype=np.uint64), np.arange(h.seq_len - 63, h.seq_len + 1, dtype=np.uint64),
                        rng.integers(0, h.seq_len + 1, 20000, dtype=np.uint64)])
    assert np.array_equal(ctx.sa_lookup(k, use_full=True), ctx.sa_lookup(k, use_full=False))
    opt = abi.default_opt()
    ga = _load("align.npz")
    seqs, off = cm.reads_for(G, 800, seed=41, indel=0.004)
    regs, n = ctx.align(opt, seqs, off)
    assert np.array_equal(n, ga["n"])
    assert np.array_equal(regs.view(np.uint8).reshape(-1, abi.ALNREG_DTYPE.itemsize), ga["regs"])
    ctx.close()
