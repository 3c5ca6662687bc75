"""Stage-seam parity of the device code against the unmodified reference CPU implementation
(SURVEY.md section 8c seams 1-5).  Bit-exact: every comparison is array equality.

Each test runs on the "emu" backend (the same kernels under the SIMT emulator, CPU, `-m "not gpu"`) and on
the "cuda" backend (`-m gpu`, through the C ABI of libbwamem_b200.so).
"""
import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import abi, lib as L

BACKENDS = [pytest.param("emu", id="emu"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]
needs_ref = pytest.mark.skipif(not cm.have_ref(), reason="oracle/_ref not built")


def _sizes(backend, emu, cuda):
    return emu if backend == "emu" else cuda


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname", ["tiny", "small"])
def test_occ_extend_sa(backend, gname):
    """bwt_occ4 / bwt_2occ4 / bwt_extend / bwt_sa (bwt.c:169,189,262,86) incl. k=-1, k=primary, k=seq_len."""
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    rng = np.random.default_rng(1)
    n = _sizes(backend, 4000, 200000)
    k = rng.integers(0, idx.seq_len + 1, size=n).astype(np.uint64)
    k[:6] = [0, idx.primary, idx.seq_len, idx.primary - 1, idx.primary + 1, 127]
    k[6] = np.uint64(2 ** 64 - 1)
    assert np.array_equal(idx.occ4(k), ctx.occ4(k))
    l = np.minimum(k + rng.integers(0, 400, size=n).astype(np.uint64), np.uint64(idx.seq_len))
    l[6] = k[6]
    a, b = idx.occ2x4(k, l), ctx.occ2x4(k, l)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    # bi-intervals: start from single-base intervals and random valid sub-intervals
    ik = np.zeros((n, 3), dtype=np.uint64)
    c = rng.integers(0, 4, size=n)
    L2 = np.array(idx.L2, dtype=np.uint64)
    ik[:, 0] = L2[c] + 1
    ik[:, 1] = L2[3 - c] + 1
    ik[:, 2] = L2[c + 1] - L2[c]
    for is_back in (0, 1):
        o0, o1 = idx.extend(ik, is_back), ctx.extend(ik, is_back)
        assert np.array_equal(o0, o1)
        # second generation: extend the children again
        ch = o0[np.arange(n), rng.integers(0, 4, size=n)]
        assert np.array_equal(idx.extend(ch, 1 - is_back), ctx.extend(ch, 1 - is_back))
    kk = k.copy()
    kk[6] = 5
    ref_sa = idx.sa_lookup(kk)
    assert np.array_equal(ref_sa, ctx.sa_lookup(kk, use_full=False))   # LF walk (bwt_sa)
    assert np.array_equal(ref_sa, ctx.sa_lookup(kk, use_full=True))    # device-expanded full SA


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_smem1a(backend):
    """bwt_smem1a (bwt.c:289) per (read, x, min_intv), incl. max_intv > 0 and ambiguous bases."""
    idx, ctx = cm.ref_index("small"), cm.context(backend, "small")
    nreads = _sizes(backend, 60, 400)
    seqs, off = cm.reads_for("small", nreads, seed=11, n_rate=0.004)
    rng = np.random.default_rng(2)
    xs = rng.integers(0, 150, size=nreads).astype(np.int32)
    mi = rng.integers(0, 4, size=nreads).astype(np.int32)
    for max_intv in (0, 20):
        out, cnt, ret = ctx.smem1a(seqs, off, xs, mi, max_intv=max_intv, cap=160)
        for r in range(nreads):
            q = seqs[off[r]:off[r + 1]]
            e, eret = idx.smem1a(q, int(xs[r]), int(mi[r]), max_intv)
            assert ret[r] == eret and cnt[r] == len(e), (r, max_intv)
            assert out[r, :cnt[r]].tobytes() == e.tobytes(), (r, max_intv)


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname", ["small", "rep"])
def test_collect_intv(backend, gname):
    """mem_collect_intv (bwamem.c:137): three passes + sort."""
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    opt = abi.default_opt()
    seqs, off = cm.reads_for(gname, _sizes(backend, 1500, 50000), seed=21, n_rate=0.001)
    iv0, c0 = idx.collect_intv(opt, seqs, off)
    iv1, c1 = ctx.collect_intv(opt, seqs, off)
    assert np.array_equal(c0, c1)
    assert iv0.tobytes() == iv1.tobytes()


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname", ["small", "rep"])
@pytest.mark.parametrize("stage", [0, 1])
def test_chains(backend, gname, stage):
    """mem_chain (bwamem.c:258) and mem_chain_flt (bwamem.c:334): chain order, seeds, w, kept, first, frac_rep."""
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    opt = abi.default_opt()
    seqs, off = cm.reads_for(gname, _sizes(backend, 1200, 40000), seed=31)
    a = idx.chains(opt, seqs, off, stage=stage, cap_chn_per_read=256, cap_seed_per_read=4096)
    b = ctx.chains(opt, seqs, off, stage=stage, cap_chn_per_read=256, cap_seed_per_read=4096)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def _ksw_problems(rng, n, maxlen=200):
    qs, ts, ql, tl = [], [], [], []
    for _ in range(n):
        lq = int(rng.integers(1, maxlen))
        q = rng.integers(0, 4, size=lq, dtype=np.uint8)
        mode = rng.integers(0, 4)
        if mode == 0:
            t = rng.integers(0, 4, size=int(rng.integers(0, maxlen)), dtype=np.uint8)
        else:
            t = list(q)
            out = []
            for b in t:
                r = rng.random()
                if r < 0.04:
                    out.append(int(rng.integers(0, 4)))
                elif r < 0.06:
                    continue
                elif r < 0.08:
                    out.extend([b, int(rng.integers(0, 4))])
                else:
                    out.append(b)
            out.extend(rng.integers(0, 4, size=int(rng.integers(0, 60))).tolist())
            t = np.array(out, dtype=np.uint8)
            if mode == 3 and len(t) > 10:
                t = t[: int(rng.integers(1, len(t)))]
        if rng.random() < 0.1 and lq > 3:
            q[int(rng.integers(0, lq))] = 4
        qs.append(q); ts.append(t); ql.append(len(q)); tl.append(len(t))
    qoff = np.concatenate([[0], np.cumsum(ql)[:-1]]).astype(np.int64)
    toff = np.concatenate([[0], np.cumsum(tl)[:-1]]).astype(np.int64)
    q = np.concatenate(qs + [np.zeros(1, np.uint8)]).astype(np.uint8)
    t = np.concatenate(ts + [np.zeros(1, np.uint8)]).astype(np.uint8)
    return (np.array(ql, np.int32), qoff, np.ascontiguousarray(q), np.array(tl, np.int32), toff, np.ascontiguousarray(t))


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_ksw_extend2_random(backend):
    """ksw_extend2 (ksw.c:380): score, qle, tle, gtle, gscore, max_off on random / mutated / truncated pairs,
    narrow and wide bands, with and without z-drop, empty targets."""
    ctx = cm.context(backend, "tiny")
    rl = cm.reflib()
    rng = np.random.default_rng(5)
    n = _sizes(backend, 250, 20000)
    for (zdrop, end_bonus, a, b, o, e) in [(100, 5, 1, 4, 6, 1), (0, 0, 1, 4, 6, 1), (20, 10, 2, 3, 4, 2), (100, 5, 1, 1, 1, 1)]:
        opt = abi.default_opt()
        opt.a, opt.b, opt.o_del, opt.o_ins, opt.e_del, opt.e_ins = a, b, o, o + 1, e, e
        abi.fill_scmat(opt)
        ql, qoff, q, tl, toff, t = _ksw_problems(rng, n)
        w = rng.choice([1, 3, 10, 50, 100, 200], size=n).astype(np.int32)
        h0 = rng.integers(1, 160, size=n).astype(np.int32)
        exp = rl.ksw_extend2(ql, qoff, q, tl, toff, t, opt, w, end_bonus, zdrop, h0)
        got = ctx.ksw_extend2(ql, qoff, q, tl, toff, t, opt, w, end_bonus, zdrop, h0)
        bad = np.nonzero((exp != got).any(axis=1))[0]
        assert len(bad) == 0, (bad[:5], exp[bad[:5]], got[bad[:5]])


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_ksw_align2_random(backend):
    """ksw_align2 (ksw.c:343): all 7 kswr_t fields; byte and 16-bit score paths, 2nd-best threshold (KSW_XSUBO),
    start coordinates (KSW_XSTART reverse pass with KSW_XSTOP), byte saturation at 255, empty targets."""
    ctx = cm.context(backend, "tiny")
    rl = cm.reflib()
    rng = np.random.default_rng(11)
    n = _sizes(backend, 300, 20000)
    for (a, b, o, e) in [(1, 4, 6, 1), (2, 3, 4, 2), (1, 1, 1, 1)]:
        opt = abi.default_opt()
        opt.a, opt.b, opt.o_del, opt.o_ins, opt.e_del, opt.e_ins = a, b, o, o + 1, e, e
        abi.fill_scmat(opt)
        ql, qoff, q, tl, toff, t = cm.sw_problems(rng, n)
        flags = rng.choice([0, cm.KSW_XSUBO, cm.KSW_XSTART, cm.KSW_XSUBO | cm.KSW_XSTART], size=n)
        byte = np.where(ql * a < 250, rng.choice([0, cm.KSW_XBYTE], size=n), 0)      # bwamem_pair.c:153
        xtra = (flags | byte | rng.integers(1, 40, size=n)).astype(np.int32)
        exp = rl.ksw_align2(ql, qoff, q, tl, toff, t, opt, xtra)
        got = ctx.ksw_align2(ql, qoff, q, tl, toff, t, opt, xtra)
        bad = np.nonzero((exp != got).any(axis=1))[0]
        assert len(bad) == 0, (bad[:5], xtra[bad[:5]], exp[bad[:5]], got[bad[:5]])
    # byte scores running into saturation (no reverse pass: the reference reads H0[-1] there)
    opt = abi.default_opt()
    ql, qoff, q, tl, toff, t = cm.sw_problems(rng, max(n // 4, 50), maxq=300, sat=True)
    xtra = np.full(len(ql), cm.KSW_XBYTE | cm.KSW_XSUBO | 19, np.int32)
    exp = rl.ksw_align2(ql, qoff, q, tl, toff, t, opt, xtra)
    got = ctx.ksw_align2(ql, qoff, q, tl, toff, t, opt, xtra)
    assert (exp[:, 0] == 255).any()
    assert np.array_equal(exp, got)


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname", ["small", "rep"])
def test_align_regs(backend, gname):
    """mem_align1_core (bwamem.c:1061) for a batch: all 88 bytes of every mem_alnreg_t."""
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    opt = abi.default_opt()
    seqs, off = cm.reads_for(gname, _sizes(backend, 400, 60000), seed=41, indel=0.004)
    r0, n0 = idx.align(opt, seqs, off, n_threads=8, cap_per_read=256)
    r1, n1 = ctx.align(opt, seqs, off)
    assert np.array_equal(n0, n1)
    assert r0.tobytes() == r1.tobytes()


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_align_edge_cases(backend):
    """Empty / too-short / all-N / N-interrupted / long / reverse / random / chimeric reads; ASCII input."""
    idx, ctx = cm.ref_index("small"), cm.context(backend, "small")
    opt = abi.default_opt()
    seqs, off = cm.ragged_reads("small")
    r0, n0 = idx.align(opt, seqs, off, cap_per_read=256)
    r1, n1 = ctx.align(opt, seqs, off)
    assert np.array_equal(n0, n1), (n0, n1)
    assert r0.tobytes() == r1.tobytes()
    # same reads as upper/lower-case ASCII
    from bwa_mem_gpu_b200 import synth
    asc = synth.to_ascii(seqs).copy()
    asc[::3] = np.char.lower(asc[::3].view("S1")).view(np.uint8)
    r2, n2 = ctx.align(opt, np.ascontiguousarray(asc), off)
    assert np.array_equal(n0, n2) and r0.tobytes() == r2.tobytes()


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_align_nondefault_options(backend):
    """Scoring / band / seed options away from the defaults (-A2 style scaling, small w, -k 15, -r 1.2, -c 50)."""
    idx, ctx = cm.ref_index("rep"), cm.context(backend, "rep")
    seqs, off = cm.reads_for("rep", _sizes(backend, 200, 20000), seed=51, sub=0.03, indel=0.006)
    for kw in (dict(a=2, b=8, o_del=12, o_ins=12, e_del=2, e_ins=2, zdrop=200, pen_clip5=10, pen_clip3=10, T=60),
               dict(w=8, min_seed_len=15, split_factor=1.2, max_occ=50, zdrop=30),
               dict(max_mem_intv=0, min_chain_weight=25, drop_ratio=0.3, mask_level=0.3, max_chain_extend=2)):
        opt = abi.default_opt()
        for k, v in kw.items():
            setattr(opt, k, v)
        abi.fill_scmat(opt)
        r0, n0 = idx.align(opt, seqs, off, cap_per_read=512)
        r1, n1 = ctx.align(opt, seqs, off)
        assert np.array_equal(n0, n1), kw
        assert r0.tobytes() == r1.tobytes(), kw


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_collect_intv_list_spill(backend):
    """Seeding kernel with a tiny shared-memory list: most reads take the queue + general-kernel path; same result."""
    idx, ctx = cm.ref_index("rep"), cm.context(backend, "rep")
    opt = abi.default_opt()
    seqs, off = cm.reads_for("rep", _sizes(backend, 600, 20000), seed=77, n_rate=0.001)
    iv0, c0 = idx.collect_intv(opt, seqs, off)
    # smem_path 0: ring-list kernel (smem3.cu, lcap 8 / 12 / 16 with spill to global); 2: fixed-list kernel with the
    # retry queue (smem2.cu); 1: general kernel only (smem.cu)
    for path, lcap in ((0, 8), (0, 12), (0, 16), (2, 2), (2, 9), (1, 0)):
        ctx.tune("smem_path", path)
        ctx.tune("smem_lcap", lcap)
        try:
            iv1, c1 = ctx.collect_intv(opt, seqs, off)
        finally:
            ctx.tune("smem_lcap", 0)
            ctx.tune("smem_path", 0)
        assert np.array_equal(c0, c1), (path, lcap)
        assert iv0.tobytes() == iv1.tobytes(), (path, lcap)


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_align_general_extension_kernel(backend):
    """The warp-per-read extension kernel (long reads / wide scores) forced onto short reads: same regions as the
    round-based fast path and as the reference."""
    idx, ctx = cm.ref_index("rep"), cm.context(backend, "rep")
    opt = abi.default_opt()
    seqs, off = cm.reads_for("rep", _sizes(backend, 300, 30000), seed=61, sub=0.02, indel=0.005)
    r0, n0 = idx.align(opt, seqs, off, n_threads=8, cap_per_read=256)
    ctx.tune("ext_path", 1)
    try:
        r1, n1 = ctx.align(opt, seqs, off)
    finally:
        ctx.tune("ext_path", 0)
    r2, n2 = ctx.align(opt, seqs, off)
    for tail in (0, 40, 1 << 30):                    # rounds only / rounds, then one tail launch once <= 40 jobs remain / tail from round 0
        ctx.tune("ext_tail", tail)
        try:
            r6, n6 = ctx.align(opt, seqs, off)
        finally:
            ctx.tune("ext_tail", 16384)
        assert np.array_equal(n0, n6) and r0.tobytes() == r6.tobytes(), tail
    ctx.tune("ext_tail", 0)                          # the scheduling variants below are those of the per-round kernels
    ctx.tune("ext_rowsync", 0)                       # byte-score kernel as a one-slot-per-iteration state machine
    try:
        r3, n3 = ctx.align(opt, seqs, off)
    finally:
        ctx.tune("ext_rowsync", 1)
    ctx.tune("ext_path", 2)                          # the 32-bit-score round kernel (ext_dp_kernel), each scheduling variant
    try:
        for tune in ({}, {"ext_rowsync": 0}, {"ext_phased": 0}, {"ext_sort": 0}):
            for k, v in tune.items():
                ctx.tune(k, v)
            try:
                r5, n5 = ctx.align(opt, seqs, off)
            finally:
                for k in tune:
                    ctx.tune(k, 1)
            assert np.array_equal(n0, n5) and r0.tobytes() == r5.tobytes(), tune
    finally:
        ctx.tune("ext_path", 0)
    for key in ("ext_phased", "ext_sort"):           # without warp-wide phases / without the job order
        ctx.tune(key, 0)
        try:
            r4, n4 = ctx.align(opt, seqs, off)
        finally:
            ctx.tune(key, 1)
        assert np.array_equal(n0, n4) and r0.tobytes() == r4.tobytes()
    ctx.tune("ext_tail", 16384)
    assert np.array_equal(n0, n1) and r0.tobytes() == r1.tobytes()
    assert np.array_equal(n0, n2) and r0.tobytes() == r2.tobytes()
    assert np.array_equal(n0, n3) and r0.tobytes() == r3.tobytes()


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname", ["small", "rep"])
def test_mate_rescue(backend, gname):
    """The mate-rescue block of mem_sam_pe (bwamem_pair.c:291-302; mem_matesw :137-206) for every pair of a batch with
    damaged mates: all 88 bytes of every region after rescue, and the number of ksw_align2 calls."""
    from bwa_mem_gpu_b200 import synth
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    opt = abi.default_opt()
    opt.flag |= abi.MEM_F_PE
    npairs = _sizes(backend, 250, 20000)
    reads, _ = synth.simulate_pairs(cm.genome(gname), npairs, seed=19, damaged_frac=0.3, damage="periodic")
    seqs = np.ascontiguousarray(reads.reshape(-1))
    off = np.arange(0, 2 * npairs + 1, dtype=np.int64) * reads.shape[1]
    regs, cnt = idx.align(opt, seqs, off, n_threads=8, cap_per_read=256)
    pes = cm.reflib_pestat(idx, opt, regs, cnt)
    assert not pes["failed"][1]                                   # FR is usable
    e_regs, e_cnt, e_sw = idx.mate_rescue(opt, pes, seqs, off, regs, cnt)
    assert e_sw > npairs // 20                                    # the case really exercises ksw_align2
    ctx.batch_upload(seqs, off)
    ctx.batch_set_regs(regs, cnt)
    n_sw, cells, _ = ctx.batch_rescue(opt, pes)
    g_regs, g_cnt = ctx.batch_download(2 * npairs, len(e_regs) + 64)
    assert n_sw == e_sw and cells > 0
    assert np.array_equal(g_cnt, e_cnt)
    assert g_regs.tobytes() == e_regs.tobytes()
    # all four orientations usable and a wide window: several SW per anchor, insertions in the middle of the list
    pes2 = pes.copy()
    pes2["failed"] = 0
    pes2["low"], pes2["high"] = 1, 900
    e_regs, e_cnt, e_sw = idx.mate_rescue(opt, pes2, seqs, off, regs, cnt)
    ctx.batch_set_regs(regs, cnt)
    n_sw, _, _ = ctx.batch_rescue(opt, pes2)
    g_regs, g_cnt = ctx.batch_download(2 * npairs, len(e_regs) + 64)
    assert n_sw == e_sw
    assert np.array_equal(g_cnt, e_cnt) and g_regs.tobytes() == e_regs.tobytes()


def _long_reads(gname, n, seed, lo=800, hi=3000, sub=0.06, indel=0.01):
    """Long noisy reads: seeds are short, so mem_flt_chained_seeds (l_query >~ 735) has real work."""
    from bwa_mem_gpu_b200 import synth
    rng = np.random.default_rng(seed)
    contigs = cm.genome(gname)
    seqs, off = [], [0]
    for i in range(n):
        l = int(rng.integers(lo, hi))
        r, _ = synth.simulate_reads(contigs, 1, read_len=l, seed=seed * 1000 + i, sub_rate=sub, indel_rate=indel)
        s = r.reshape(-1)
        if i % 7 == 3:                                     # chimeric: second half from elsewhere
            r2, _ = synth.simulate_reads(contigs, 1, read_len=l, seed=seed * 1000 + i + 500000, sub_rate=sub, indel_rate=indel)
            s = np.concatenate([s[: l // 2], r2.reshape(-1)[l // 2:]])
        seqs.append(s); off.append(off[-1] + len(s))
    return np.ascontiguousarray(np.concatenate(seqs).astype(np.uint8)), np.array(off, dtype=np.int64)


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname", ["small", "rep"])
@pytest.mark.parametrize("scoring", ["default", "ont"])
def test_flt_chained_seeds_long_reads(backend, gname, scoring):
    """mem_flt_chained_seeds / mem_seed_sw (bwamem.c:605-622, 578-603): surviving seeds and their SW scores, from the
    one-alignment-per-lane kernel (seedsw2.cu, default) and from the lane-group restatement of the striped code."""
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    opt = abi.default_opt()
    if scoring == "ont":                                   # cheap mismatches, cheap and unequal gap costs (in the spirit of -x ont2d)
        opt.a, opt.b, opt.o_del, opt.o_ins, opt.e_del, opt.e_ins = 2, 3, 2, 7, 1, 2
        opt.min_seed_len = 14
        abi.fill_scmat(opt)
    seqs, off = _long_reads(gname, _sizes(backend, 24, 600), seed=3)
    a1 = idx.chains(opt, seqs, off, stage=1, cap_chn_per_read=512, cap_seed_per_read=16384)
    a = idx.chains(opt, seqs, off, stage=2, cap_chn_per_read=512, cap_seed_per_read=16384)
    b = ctx.chains(opt, seqs, off, stage=2, cap_chn_per_read=512, cap_seed_per_read=16384)
    ctx.tune("ssw_path", 1)
    try:
        c = ctx.chains(opt, seqs, off, stage=2, cap_chn_per_read=512, cap_seed_per_read=16384)
    finally:
        ctx.tune("ssw_path", 0)
    if scoring == "default":
        assert len(a[2]) < len(a1[2])                      # the filter really removed seeds
    for x, y, z in zip(a, b, c):
        assert np.array_equal(x, y) and np.array_equal(x, z)


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_align_long_reads(backend):
    """mem_align1_core on 0.8-3 kbp noisy reads (seed SW filter + wide-band extension + region patching)."""
    idx, ctx = cm.ref_index("small"), cm.context(backend, "small")
    opt = abi.default_opt()
    seqs, off = _long_reads("small", _sizes(backend, 16, 400), seed=5)
    r0, n0 = idx.align(opt, seqs, off, n_threads=8, cap_per_read=512)
    r1, n1 = ctx.align(opt, seqs, off)
    assert np.array_equal(n0, n1)
    assert r0.tobytes() == r1.tobytes()


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("tune", [{"force_wide": 1}, {"force_wide": 1, "kmer_k": 0}, {"force_wide": 2}, {"force_wide": 2, "kmer_k": 0, "smem_lcap": 8},
                                  {"kmer_k": 0}, {"kmer_k": 6}, {"kmer_k": 12},
                                  {"smem_path": 3}, {"smem_path": 3, "force_wide": 1}, {"smem_path": 3, "force_wide": 2}, {"smem_path": 3, "split_p3": 0}])
def test_seeding_variants(backend, tune):
    """The 64-bit-position templates of the seeding kernel (used when the doubled text reaches 2^32 rows) forced on a
    small genome, k-mer interval tables of other depths, and the long-read variant that fetches the read's codes from
    global memory (smem_path 3): intervals and regions unchanged."""
    from bwa_mem_gpu_b200 import lib as L
    gname = "rep"
    lib = cm.backend_lib(backend)
    ctx = L.Context(0, lib)
    for k, v in tune.items():
        ctx.tune(k, v)
    ctx.upload_host_index(cm.context(backend, gname)._host_index)
    idx = cm.ref_index(gname)
    opt = abi.default_opt()
    seqs, off = cm.reads_for(gname, _sizes(backend, 300, 20000), seed=57, indel=0.004, n_rate=0.002)
    a = idx.collect_intv(opt, seqs, off)
    b = ctx.collect_intv(opt, seqs, off)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    r0, n0 = idx.align(opt, seqs, off, n_threads=8, cap_per_read=256)
    r1, n1 = ctx.align(opt, seqs, off)
    assert np.array_equal(n0, n1) and r0.tobytes() == r1.tobytes()


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("min_seed_len", [19, 13, 10])
def test_seeding_table_counters(backend, min_seed_len):
    """The k-mer table path (and its bit-only list entries, used while min_seed_len exceeds the table depth) answers the
    same bwt_extend calls as the plain path: same intervals, and the same extension / two-block counts (the roofline's
    algorithmic traffic must not depend on how an extension was served)."""
    from bwa_mem_gpu_b200 import lib as L
    gname = "rep"
    lib = cm.backend_lib(backend)
    idx = cm.ref_index(gname)
    opt = abi.default_opt()
    opt.min_seed_len = min_seed_len
    seqs, off = cm.reads_for(gname, _sizes(backend, 400, 20000), seed=58, indel=0.004, n_rate=0.002)
    a = idx.collect_intv(opt, seqs, off)
    got = []
    for k, stats in ((0, 1), (12, 1), (12, 0)):              # stats off: the production kernels (entries settled in groups)
        ctx = L.Context(0, lib)
        ctx.tune("kmer_k", k)
        ctx.tune("seed_stats", stats)
        ctx.upload_host_index(cm.context(backend, gname)._host_index)
        b = ctx.collect_intv(opt, seqs, off)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
        if stats:
            ctx.align(opt, seqs, off)
            c = ctx.counters()
            got.append((c["extend"], c["two_block"]))
        ctx.close()
    assert got[0] == got[1] and got[0][0] > 0


def _ragged_noisy_reads(gname, n, seed):
    """Reads of 12-170 bases cut from the genome (both strands), 2 % substitutions and 3 % N: sweeps that start, stop and
    restart inside the k-mer table's range in every combination."""
    rng = np.random.default_rng(seed)
    g = np.concatenate(cm.genome(gname))
    parts = []
    for _ in range(n):
        ln = int(rng.integers(12, 171))
        p = int(rng.integers(0, len(g) - ln))
        r = g[p: p + ln].copy()
        if rng.random() < 0.5:
            r = (3 - r)[::-1].copy()
        m = rng.random(ln) < 0.02
        r[m] = (r[m] + rng.integers(1, 4, int(m.sum()))) % 4
        r[rng.random(ln) < 0.03] = 4
        parts.append(r.astype(np.uint8))
    seqs = np.ascontiguousarray(np.concatenate(parts))
    off = np.concatenate([[0], np.cumsum([len(x) for x in parts])]).astype(np.int64)
    return seqs, off


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("optset", [dict(), dict(min_seed_len=13), dict(min_seed_len=13, split_factor=1.0, split_width=30),
                                    dict(min_seed_len=14, max_mem_intv=0), dict(min_seed_len=15, max_mem_intv=200, split_width=2)])
def test_seeding_ragged_reads_table_depths(backend, optset):
    """mem_collect_intv on short, N-rich, ragged reads with table depths below, at and above typical sweep lengths and
    re-seeding / third-pass options moved around: the production kernels (bit entries, group decisions) against the
    reference, interval for interval."""
    from bwa_mem_gpu_b200 import lib as L
    gname = "rep"
    lib = cm.backend_lib(backend)
    idx = cm.ref_index(gname)
    opt = abi.default_opt()
    for k, v in optset.items():
        setattr(opt, k, v)
    seqs, off = _ragged_noisy_reads(gname, _sizes(backend, 500, 30000), seed=63)
    a = idx.collect_intv(opt, seqs, off)
    for depth in (5, 9, 12):
        ctx = L.Context(0, lib)
        ctx.tune("kmer_k", depth)
        ctx.upload_host_index(cm.context(backend, gname)._host_index)
        b = ctx.collect_intv(opt, seqs, off)
        ctx.close()
        for x, y in zip(a, b):
            assert np.array_equal(x, y), depth


def _check_reg2aln(idx, ctx, opt, seqs, off):
    regs, cnt = idx.align(opt, seqs, off, n_threads=8, cap_per_read=512)
    a0, c0, m0 = idx.reg2aln(opt, seqs, off, regs, cnt)
    a1, c1, m1 = ctx.reg2aln(opt, seqs, off, regs, cnt)
    assert len(a0) == len(a1) == len(regs)
    for k, name in enumerate(["pos", "rid", "is_rev", "n_cigar", "NM", "score", "sub", "l_md"]):
        assert np.array_equal(a0[:, k], a1[name].astype(np.int64)), name
    assert np.array_equal(c0, c1)                          # CIGAR words in order = identical CIGARs (n_cigar matched above)
    assert m0.tobytes() == m1.tobytes()                    # MD strings
    assert np.array_equal(a1["cigar_off"], np.concatenate([[0], np.cumsum(a1["n_cigar"])[:-1]]))
    assert np.array_equal(a1["md_off"], np.concatenate([[0], np.cumsum(a1["l_md"])[:-1]]))
    return a1, c1


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname", ["small", "rep"])
def test_reg2aln(backend, gname):
    """mem_reg2aln (bwamem.c:1099-1169): position, CIGAR, NM and MD of every region -- short reads with indels, ragged
    reads (clips, reverse strand, chimeric), non-default scoring."""
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    opt = abi.default_opt()
    seqs, off = cm.reads_for(gname, _sizes(backend, 250, 20000), seed=77, sub=0.02, indel=0.01)
    a, c = _check_reg2aln(idx, ctx, opt, seqs, off)
    ops = c & 0xf
    assert (ops == 1).any() and (ops == 2).any() and (ops == 3).any()      # insertions, deletions and clips occurred
    assert a["is_rev"].any() and not a["is_rev"].all()
    ctx.tune("r2a_small", 0)                               # full H / E row for every DP region (no ring pass)
    try:
        _check_reg2aln(idx, ctx, opt, seqs, off)
    finally:
        ctx.tune("r2a_small", 1)
    seqs, off = cm.ragged_reads(gname)
    _check_reg2aln(idx, ctx, opt, seqs, off)
    opt2 = abi.default_opt()
    opt2.a, opt2.b, opt2.o_del, opt2.o_ins, opt2.e_del, opt2.e_ins = 2, 3, 2, 7, 1, 2
    abi.fill_scmat(opt2)
    seqs, off = cm.reads_for(gname, _sizes(backend, 120, 5000), seed=78, sub=0.03, indel=0.01)
    _check_reg2aln(idx, ctx, opt2, seqs, off)


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_reg2aln_long_reads(backend):
    idx, ctx = cm.ref_index("small"), cm.context(backend, "small")
    seqs, off = _long_reads("small", _sizes(backend, 6, 200), seed=9)
    _check_reg2aln(idx, ctx, abi.default_opt(), seqs, off)


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_align_batch_pipeline_edges(backend):
    """b200_align_batch (uploads / chunks / downloads overlapped): many small chunks, offsets that do not start at 0, an
    output capacity that is too small (-needed, nothing truncated silently), an empty batch."""
    import ctypes as C
    idx, ctx = cm.ref_index("small"), cm.context(backend, "small")
    opt = abi.default_opt()
    n = _sizes(backend, 300, 20000)
    seqs, off = cm.reads_for("small", n, seed=4)
    r0, n0 = idx.align(opt, seqs, off, n_threads=8, cap_per_read=256)
    ctx.tune("chunk_reads", max(64, n // 7))
    try:
        r1, n1 = ctx.align(opt, seqs, off)
        r2, n2 = ctx.align(opt, seqs, off[n // 3:].copy())
        out = np.zeros(5, dtype=abi.ALNREG_DTYPE); cnt = np.zeros(n, dtype=np.int32)
        tot = ctx.lib.b200_align_batch(ctx.h, C.byref(opt), C.c_int(n), seqs.ctypes.data_as(C.POINTER(C.c_uint8)),
                                       off.ctypes.data_as(C.POINTER(C.c_int64)), C.c_int64(5), C.c_void_p(out.ctypes.data),
                                       cnt.ctypes.data_as(C.POINTER(C.c_int32)))
        r3, n3 = ctx.align(opt, seqs, off[:1].copy())
    finally:
        ctx.tune("chunk_reads", 1 << 20)
    assert np.array_equal(n0, n1) and r0.tobytes() == r1.tobytes()
    k = int(n0[: n // 3].sum())
    assert np.array_equal(n0[n // 3:], n2) and r0[k:].tobytes() == r2.tobytes()
    assert tot == -len(r0) and np.array_equal(cnt, n0)
    assert len(r3) == 0 and len(n3) == 0


def _nt4(seqs):
    lut = np.full(256, 4, dtype=np.uint8)
    for k, ch in enumerate("ACGT"):
        lut[ord(ch)] = lut[ord(ch.lower())] = k
    lut[:4] = np.arange(4)
    return lut[np.asarray(seqs, dtype=np.uint8)]


def _check_reg2sam_se(idx, ctx, opt, seqs, off, n_processed=0, resident=True):
    """b200_reg2sam_se_batch + b200_sam_se_format against the reference's worker2 (mem_mark_primary_se,
    mem_approx_mapq_se, mem_reg2sam): region arrays byte for byte, MAPQ, and the SAM text byte for byte."""
    regs, cnt = idx.align(opt, seqs, off, n_threads=8, cap_per_read=512)
    r0, q0, sam0 = idx.reg2sam_se(opt, _nt4(seqs), off, regs, cnt, n_processed)
    r1, sel, aln, cig, md = ctx.reg2sam_se(opt, seqs, off, regs, cnt, n_processed)
    assert r0.tobytes() == r1.tobytes()                    # order, hash, sub, sub_n, alt_sc, secondary, secondary_all
    rec = (sel["kind"] & 1) != 0
    first = np.zeros(len(regs), dtype=bool)                # the first record of a read carries mem_approx_mapq_se unchanged
    g = 0
    for c in cnt:
        ks = np.nonzero(rec[g:g + c])[0]
        if len(ks):
            first[g + ks[0]] = True
        g += c
    assert np.array_equal(sel["mapq"][first], q0[first])
    names = [f"r{n_processed + i}" for i in range(len(off) - 1)]
    sam1, soff = ctx.sam_se(opt, idx.anns()[3], names, seqs, off, cnt, sel, aln, cig, md)
    if sam0 != sam1:
        l0, l1 = sam0.split(b"\n"), sam1.split(b"\n")
        for x, y in zip(l0, l1):
            assert x == y
        assert len(l0) == len(l1)
    assert soff[-1] == len(sam1) and sam1[soff[1]:].startswith(names[1].encode()) if len(names) > 1 else True
    # the whole way on the device: worker1, worker2 and the device SAM formatter (samdev.cu) through b200_reads_to_sam
    nf, no = L.flat_strings(names)
    sam3, soff3, _ = L.reads_to_sam([ctx], opt, idx.anns()[3], seqs, off, nf, no, n_processed=n_processed, sub_reads=max(2, (len(off) - 1) // 3))
    if sam3.tobytes() != sam1:
        for x, y in zip(sam1.split(b"\n"), sam3.tobytes().split(b"\n")):
            assert x == y
    assert sam3.tobytes() == sam1 and np.array_equal(soff3, soff)
    if not resident:
        return r1, sel, aln, sam1
    # the same on the batch the context has just aligned (b200_batch_reg2sam_se: nothing uploaded), twice: the resident
    # regions are not modified by the call
    regs_d, cnt_d = ctx.align(opt, seqs, off)
    assert np.array_equal(cnt_d, cnt) and regs_d.tobytes() == regs.tobytes()
    for _ in range(2):
        r2, sel2, aln2, cig2, md2 = ctx.batch_reg2sam_se(opt, cnt_d, n_processed)
        assert r2.tobytes() == r1.tobytes() and sel2.tobytes() == sel.tobytes() and aln2.tobytes() == aln.tobytes()
        assert np.array_equal(cig2, cig) and md2.tobytes() == md.tobytes()
    # and as one pipelined call, cut into sub-batches (text of one written while the next is aligned)
    for sub in (0, max(1, (len(off) - 1) // 3)):
        sam2, soff2 = ctx.align_sam_se(opt, idx.anns()[3], names, seqs, off, n_processed, sub_reads=sub)
        assert sam2 == sam1 and np.array_equal(soff2, soff)
    return r1, sel, aln, sam1


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname", ["small", "rep", "dup"])
def test_reg2sam_se(backend, gname):
    """worker2's single-end half (bwamem.c:1195-1205): primary marking with the hash tie-break, MAPQ, record selection,
    XA, SAM text -- clean reads, repeat-rich reads (secondary hits, sub / sub_n, XA), ragged / chimeric reads
    (supplementary records, hard clips, SA tags), reads with no hit (unmapped record), -a, -Y, -5 and -M."""
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    opt = abi.default_opt()
    seqs, off = cm.reads_for(gname, _sizes(backend, 250, 20000), seed=91, sub=0.02, indel=0.01)
    r1, sel, aln, sam = _check_reg2sam_se(idx, ctx, opt, seqs, off, n_processed=12345)
    if gname == "dup":                                     # duplicated segments and an ALT contig (tests/common.py dup_genome)
        assert b"XA:Z:" in sam and b"pa:f:" in sam and (r1["sub_n"] > 0).any() and (sel["kind"] == 3).any()
        assert (r1["secondary"] == 0x7fffffff).any() and (r1["alt_sc"] > 0).any()
        assert (r1["secondary_all"] >= 0).sum() > (sel["kind"] & 2 != 0).sum()      # some reads exceed max_XA_hits: no XA for them
    seqs, off = cm.ragged_reads(gname)
    r1, sel, aln, sam = _check_reg2sam_se(idx, ctx, opt, seqs, off)
    assert b"SA:Z:" in sam and (aln["flag"] & 0x800).any()
    assert b"\t4\t*\t0\t0\t*" in sam                       # a read without any record
    for flag in (0x8, 0x200, 0x800, 0x10, 0x8 | 0x2000 | 0x1000):      # MEM_F_ALL, SOFTCLIP, PRIMARY5, NO_MULTI, ALL|XB|KEEP_SUPP_MAPQ
        o = abi.default_opt()
        o.flag |= flag
        _check_reg2sam_se(idx, ctx, o, seqs, off, n_processed=7, resident=False)
    o = abi.default_opt()
    o.mapQ_coef_len = 0                                    # the other MAPQ formula (log(seedcov))
    _check_reg2sam_se(idx, ctx, o, seqs, off, resident=False)


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_reg2sam_se_long_reads(backend):
    idx, ctx = cm.ref_index("small"), cm.context(backend, "small")
    seqs, off = _long_reads("small", _sizes(backend, 6, 200), seed=19)
    _check_reg2sam_se(idx, ctx, abi.default_opt(), seqs, off, n_processed=3, resident=backend != "emu")   # emulated long-read aligns are slow


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_interval_capacity_retry_and_error(backend):
    """A chunk whose reads need more interval slots than they were given is re-run with the reported need (same result as
    the reference); with cap_retry=0 the call fails -- in the reads -> SAM call too, which must not produce text."""
    idx, ctx = cm.ref_index("rep"), cm.context(backend, "rep")
    opt = abi.default_opt()
    seqs, off = cm.reads_for("rep", _sizes(backend, 300, 5000), seed=91)
    r0, n0 = idx.align(opt, seqs, off, cap_per_read=512)
    try:
        ctx.tune("cap_intv", 1)
        r1, n1 = ctx.align(opt, seqs, off)
        assert np.array_equal(n0, n1) and r0.tobytes() == r1.tobytes()
        ctx.tune("cap_intv", 1)
        ctx.tune("cap_retry", 0)
        with pytest.raises(L.B200Error, match="interval capacity"):
            ctx.align(opt, seqs, off)
        names = [f"r{i}" for i in range(len(off) - 1)]
        with pytest.raises(L.B200Error, match="interval capacity"):
            ctx.align_sam_se(opt, ctx._host_index.names, names, seqs, off)
    finally:
        ctx.tune("cap_intv", 0)
        ctx.tune("cap_retry", 1)
