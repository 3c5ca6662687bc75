"""Own index builders (host SA-IS; device prefix-key sort) vs the reference's `bwa index` arrays: byte-identical."""
import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import abi, lib as L

needs_ref = pytest.mark.skipif(not cm.have_ref(), reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("gname", ["tiny", "small", "rep"])
def test_host_index_matches_reference(gname):
    idx = cm.ref_index(gname)
    h = L.HostIndex(cm.genome(gname), cm.backend_lib("emu"))
    assert (h.primary, h.L2, h.seq_len, h.bwt_size, h.n_sa, h.sa_intv) == \
           (idx.primary, idx.L2, idx.seq_len, idx.bwt_size, idx.n_sa, idx.sa_intv)
    assert np.array_equal(h.bwt_words(), idx.bwt_words())
    assert np.array_equal(h.sa(), idx.sa())
    nb = idx.l_pac // 4 + (1 if idx.l_pac % 4 else 0)
    assert np.array_equal(h.pac()[:nb], idx.pac()[:nb])
    h.close()


def _device_vs_host(backend, gname, tune=()):
    lib = cm.backend_lib(backend)
    g = cm.genome(gname)
    h = L.HostIndex(g, lib)
    ctx = L.Context(0, lib)
    for k, v in tune:
        ctx.tune(k, v)
    d = ctx.build_index_device(g)
    assert (d.primary, d.L2, d.seq_len, d.bwt_size, d.n_sa, d.sa_intv) == (h.primary, h.L2, h.seq_len, h.bwt_size, h.n_sa, h.sa_intv)
    assert np.array_equal(d.bwt_words(), h.bwt_words())
    assert np.array_equal(d.sa(), h.sa())
    assert np.array_equal(d.pac(), h.pac())
    # the resident full SA and Occ structures answer like those derived from an uploaded index
    ctx2 = cm.context(backend, gname)
    rng = np.random.default_rng(11)
    k = rng.integers(0, h.seq_len + 1, 4000).astype(np.uint64)
    k[0] = 0; k[1] = h.seq_len; k[2] = h.primary
    assert np.array_equal(ctx.sa_lookup(k, use_full=True), ctx2.sa_lookup(k, use_full=True))
    assert np.array_equal(ctx.sa_lookup(k[3:300], use_full=False), ctx2.sa_lookup(k[3:300], use_full=False))
    assert np.array_equal(ctx.occ4(k), ctx2.occ4(k))
    # and the whole path gives the same regions
    seqs, off = cm.reads_for(gname, 200, seed=21)
    opt = cm.abi.default_opt()
    r1, n1 = ctx.align(opt, seqs, off)
    r2, n2 = ctx2.align(opt, seqs, off)
    assert np.array_equal(n1, n2) and r1.tobytes() == r2.tobytes()
    ctx.close(); h.close()
    return d


@pytest.mark.parametrize("gname,tune", [("tiny", ()), ("tiny", (("idx_group_cap", 3000),)), ("rep", (("idx_group_cap", 70000),)),
                                        ("small", ())])
def test_device_index_matches_host_emu(gname, tune):
    d = _device_vs_host("emu", gname, tune)
    if tune:
        assert d.sort_groups > 1
    if gname == "rep":
        assert d.host_ordered > 0                 # repeat copies collide on 32 bases: the host ordering path ran


def test_device_index_tie_cap_emu():
    lib = cm.backend_lib("emu")
    ctx = L.Context(0, lib)
    ctx.tune("idx_tie_cap", 8)
    with pytest.raises(L.B200Error, match="tie cap"):
        ctx.build_index_device(cm.genome("rep"))
    ctx.close()


def test_pack_pac_ragged():
    rng = np.random.default_rng(3)
    contigs = [rng.integers(0, 4, n).astype(np.uint8) for n in (1, 2, 3, 5, 9, 4, 1, 1, 1, 1, 6, 0, 7)]
    pac, l_pac = L.pack_pac(contigs)
    fwd = np.concatenate(contigs)
    want = np.zeros(l_pac // 4 + 1, dtype=np.uint8)
    for i, c in enumerate(fwd):
        want[i >> 2] |= int(c) << ((~i & 3) << 1)
    assert l_pac == len(fwd) and np.array_equal(pac, want)


@pytest.mark.gpu
@pytest.mark.parametrize("gname,tune", [("tiny", (("idx_group_cap", 3000),)), ("rep", ()), ("small", ())])
def test_device_index_matches_host_cuda(gname, tune):
    _device_vs_host("cuda", gname, tune)


@pytest.mark.gpu
@pytest.mark.parametrize("with_sa", [False, True])
def test_index_save_load_replicate(tmp_path, with_sa):
    """On-disk device-ready index: save from one context, load into another (with and without the full SA in the file);
    a third context borrows the loaded index (b200_index_replicate); all give the regions of the original."""
    gname = "rep"
    ctx = cm.context("cuda", gname)
    opt = abi.default_opt()
    seqs, off = cm.reads_for(gname, 5000, seed=61, indel=0.004)
    r0, n0 = ctx.align(opt, seqs, off)
    path = str(tmp_path / "idx.b200")
    ctx.save_index(path, with_full_sa=with_sa)
    c2 = L.Context(0, ctx.lib)
    c2.load_index(path)
    r1, n1 = c2.align(opt, seqs, off)
    assert np.array_equal(n0, n1) and r0.tobytes() == r1.tobytes()
    assert np.array_equal(c2.index_info()[:11], ctx.index_info()[:11])
    c3 = L.Context(0, ctx.lib)
    c3.replicate_index_from(c2)
    r2, n2 = c3.align(opt, seqs, off)
    assert np.array_equal(n0, n2) and r0.tobytes() == r2.tobytes()
    c3.close(); c2.close()
    with pytest.raises(L.B200Error):
        L.Context(0, ctx.lib).load_index(str(tmp_path / "nope"))
