"""The scalar drop-in symbols (include/bwamem_b200.h, BWAMEM_B200_NO_DROPIN block): same names, signatures and side
effects as the reference's bwt.h / ksw.h / bwamem.c functions, called with the REFERENCE's own bwt_t / bntseq_t
structs (as a host program that links libbwamem_b200 in place of bwt.o / ksw.o would)."""
import ctypes as C

import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import abi

BACKENDS = [pytest.param("emu", id="emu"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]
needs_ref = pytest.mark.skipif(not cm.have_ref(), reason="oracle/_ref not built")


class KswR(C.Structure):          # kswr_t, ksw.h:14-19
    _fields_ = [(k, C.c_int) for k in ("score", "te", "qe", "score2", "te2", "tb", "qb")]


class IntvV(C.Structure):         # bwtintv_v, bwt.h:66 (this fork: int n, m)
    _fields_ = [("n", C.c_int), ("m", C.c_int), ("a", C.c_void_p)]


def _u8p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_scalar_fmindex_symbols(backend):
    """bwt_occ4 / bwt_2occ4 / bwt_extend / bwt_sa (bwt.h:103-117) on the reference's bwt_t."""
    idx = cm.ref_index("tiny")
    lib = cm.backend_lib(backend)
    bwt = C.c_void_p(idx.bwt_struct())
    rng = np.random.default_rng(2)
    lib.bwt_sa.restype = C.c_uint64
    for _ in range(12):
        k, l = sorted(int(x) for x in rng.integers(0, idx.seq_len + 1, size=2))
        ck, cl = (C.c_uint64 * 4)(), (C.c_uint64 * 4)()
        lib.bwt_occ4(bwt, C.c_uint64(k), ck)
        assert list(ck) == list(idx.occ4(np.array([k], np.uint64))[0])
        lib.bwt_2occ4(bwt, C.c_uint64(k), C.c_uint64(l), ck, cl)
        ek, el = idx.occ2x4(np.array([k], np.uint64), np.array([l], np.uint64))
        assert list(ck) == list(ek[0]) and list(cl) == list(el[0])
        assert lib.bwt_sa(bwt, C.c_uint64(k)) == int(idx.sa_lookup(np.array([k], np.uint64))[0])
    # bwt_extend from the single-base intervals, both directions
    for c in range(4):
        ik = np.array([[idx.L2[c] + 1, idx.L2[3 - c] + 1, idx.L2[c + 1] - idx.L2[c]]], dtype=np.uint64)
        for is_back in (0, 1):
            ok = (C.c_uint64 * 16)()                      # 4 x bwtintv_t {x[3], info}
            iv = (C.c_uint64 * 4)(*[int(v) for v in ik[0]], 0)
            lib.bwt_extend(bwt, iv, ok, C.c_int(is_back))
            exp = idx.extend(ik, is_back)[0]
            got = np.array(list(ok), dtype=np.uint64).reshape(4, 4)[:, :3]
            assert np.array_equal(got, np.asarray(exp).reshape(4, 3))


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_scalar_ksw_symbols(backend):
    """ksw_extend2 (ksw.h:108) with nullable out-pointers; ksw_align2 (ksw.h:64) incl. the kswq_t** profile cache."""
    lib = cm.backend_lib(backend)
    rl = cm.reflib()
    opt = abi.default_opt()
    rng = np.random.default_rng(9)
    ql, qoff, q, tl, toff, t = cm.sw_problems(rng, 12)
    xtra = np.full(12, cm.KSW_XSUBO | cm.KSW_XSTART | 19, np.int32)
    xtra[::2] |= cm.KSW_XBYTE
    exp = rl.ksw_align2(ql, qoff, q, tl, toff, t, opt, xtra)
    lib.ksw_align2.restype = KswR
    for i in range(12):
        qi, ti = q[qoff[i]:qoff[i] + ql[i]].copy(), t[toff[i]:toff[i] + tl[i]].copy()
        qry = C.c_void_p(None)
        for rep in range(2):                              # 2nd call reuses the cached profile; width comes from it
            x = int(xtra[i]) if rep == 0 else int(xtra[i]) ^ cm.KSW_XBYTE
            r = lib.ksw_align2(int(ql[i]), _u8p(qi), int(tl[i]), _u8p(ti), 5, opt.mat, opt.o_del, opt.e_del, opt.o_ins,
                               opt.e_ins, x, C.byref(qry))
            assert [r.score, r.te, r.qe, r.score2, r.te2, r.tb, r.qb] == list(exp[i])
            assert qry.value
        C.CDLL(None).free(qry)
        r = lib.ksw_align2(int(ql[i]), _u8p(qi), int(tl[i]), _u8p(ti), 5, opt.mat, opt.o_del, opt.e_del, opt.o_ins,
                           opt.e_ins, int(xtra[i]), None)
        assert [r.score, r.te, r.qe, r.score2, r.te2, r.tb, r.qb] == list(exp[i])
    w = np.full(12, 100, np.int32); h0 = np.full(12, 30, np.int32)
    exp = rl.ksw_extend2(ql, qoff, q, tl, toff, t, opt, w, 5, 100, h0)
    for i in range(12):
        qi, ti = q[qoff[i]:qoff[i] + ql[i]].copy(), t[toff[i]:toff[i] + tl[i]].copy()
        o = [C.c_int() for _ in range(5)]
        sc = lib.ksw_extend2(int(ql[i]), _u8p(qi), int(tl[i]), _u8p(ti), 5, opt.mat, opt.o_del, opt.e_del, opt.o_ins,
                             opt.e_ins, 100, 5, 100, 30, *[C.byref(x) for x in o])
        assert [sc] + [x.value for x in o] == list(exp[i])
        assert lib.ksw_extend2(int(ql[i]), _u8p(qi), int(tl[i]), _u8p(ti), 5, opt.mat, opt.o_del, opt.e_del, opt.o_ins,
                               opt.e_ins, 100, 5, 100, 30, None, None, None, None, None) == exp[i][0]


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_scalar_smem1a_and_align1(backend):
    """bwt_smem1a (bwt.h:124) and mem_align1_core (bwamem.c:1061) one read at a time, on the reference's structs."""
    idx = cm.ref_index("tiny")
    lib = cm.backend_lib(backend)
    bwt, bns = C.c_void_p(idx.bwt_struct()), C.c_void_p(idx.bns_struct())
    pac = C.c_void_p(idx.pac().ctypes.data)
    opt = abi.default_opt()
    seqs, off = cm.reads_for("tiny", 6, seed=77, read_len=120, indel=0.01)
    regs, nreg = idx.align(opt, seqs, off, cap_per_read=64)

    class AlnV(C.Structure):
        _fields_ = [("n", C.c_int), ("m", C.c_int), ("a", C.c_void_p)]          # bwa.h:165
    lib.mem_align1_core.restype = AlnV
    k = 0
    for i in range(6):
        s = np.ascontiguousarray(seqs[off[i]:off[i + 1]]).copy()
        for x in (0, 37, 90):
            mem = IntvV(0, 0, None)
            ret = lib.bwt_smem1a(bwt, len(s), _u8p(s), x, 1, C.c_uint64(0), C.byref(mem), None)
            e_iv, e_ret = idx.smem1a(s, x)
            got = np.ctypeslib.as_array(C.cast(mem.a, C.POINTER(C.c_uint64)), shape=(mem.n, 4)).copy() if mem.n else np.zeros((0, 4), np.uint64)
            assert ret == e_ret and np.array_equal(got, e_iv.view(np.uint64).reshape(-1, 4))
            C.CDLL(None).free(C.c_void_p(mem.a))
        v = lib.mem_align1_core(C.byref(opt), bwt, bns, pac, len(s), s.ctypes.data_as(C.c_char_p), None)
        assert v.n == nreg[i]
        got = bytes((C.c_char * (v.n * abi.ALNREG_DTYPE.itemsize)).from_address(v.a)) if v.n else b""
        assert got == regs[k:k + nreg[i]].tobytes()
        k += int(nreg[i])
        C.CDLL(None).free(C.c_void_p(v.a))


@needs_ref
@pytest.mark.gpu
def test_scalar_shims_from_several_threads():
    """The reference's primitives are re-entrant and called from kt_for workers; the same-name shims serialise per context."""
    import threading
    rl = cm.reflib()
    lib = cm.backend_lib("cuda")
    opt = abi.default_opt()
    rng = np.random.default_rng(5)
    n = 64
    ql = rng.integers(20, 120, size=n).astype(np.int32); tl = (ql + rng.integers(0, 40, size=n)).astype(np.int32)
    qoff = np.concatenate([[0], np.cumsum(ql)[:-1]]).astype(np.int64); toff = np.concatenate([[0], np.cumsum(tl)[:-1]]).astype(np.int64)
    q = rng.integers(0, 4, size=int(ql.sum()), dtype=np.uint8); t = rng.integers(0, 4, size=int(tl.sum()), dtype=np.uint8)
    for i in range(n):
        m = min(ql[i], tl[i]); t[toff[i]:toff[i] + m] = q[qoff[i]:qoff[i] + m]
    t[rng.random(len(t)) < 0.05] = 0
    w = np.full(n, 50, np.int32); h0 = rng.integers(10, 60, size=n).astype(np.int32)
    exp = rl.ksw_extend2(ql, qoff, q, tl, toff, t, opt, w, 5, 100, h0)
    got = np.zeros(n, dtype=np.int64)

    def work(lo, hi):
        for i in range(lo, hi):
            qi = np.ascontiguousarray(q[qoff[i]:qoff[i] + ql[i]]); ti = np.ascontiguousarray(t[toff[i]:toff[i] + tl[i]])
            got[i] = lib.ksw_extend2(int(ql[i]), _u8p(qi), int(tl[i]), _u8p(ti), 5, opt.mat, opt.o_del, opt.e_del, opt.o_ins,
                                     opt.e_ins, int(w[i]), 5, 100, int(h0[i]), None, None, None, None, None)
    th = [threading.Thread(target=work, args=(k * 16, k * 16 + 16)) for k in range(4)]
    [x.start() for x in th]; [x.join() for x in th]
    assert np.array_equal(got, exp[:, 0])
