"""b200_ksw_extend2_noalloc (include/bwamem_b200.h): the allocation-free variant of b200_ksw_extend2_batch, checked
against that entry point (bit-exact, the tolerance of every test here) on two calls in a row -- the second call finds the
first call's rows and work counter in the shared static scratch array.  Collected last on purpose;
passes alone in a fresh process and after the other tests.  Device memory for the caller's side comes from
torch (plumbing only)."""
import ctypes as C

import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import abi

pytestmark = pytest.mark.gpu


def _problems(seed, n, max_q):
    rng = np.random.default_rng(seed)
    ql = rng.integers(1, max_q + 1, size=n).astype(np.int32)
    tl = (ql + rng.integers(-20, 60, size=n)).clip(1, None).astype(np.int32)
    qoff = np.concatenate([[0], np.cumsum(ql)[:-1]]).astype(np.int64)
    toff = np.concatenate([[0], np.cumsum(tl)[:-1]]).astype(np.int64)
    q = rng.integers(0, 4, size=int(ql.sum()), dtype=np.uint8)
    t = rng.integers(0, 4, size=int(tl.sum()), dtype=np.uint8)
    for i in range(n):                                   # targets = mutated copies of the queries, so the DP goes somewhere
        m = min(ql[i], tl[i])
        seg = q[qoff[i]:qoff[i] + m].copy()
        mut = rng.random(m) < 0.06
        seg[mut] = (seg[mut] + rng.integers(1, 4, size=int(mut.sum()))) % 4
        t[toff[i]:toff[i] + m] = seg
    q[rng.random(len(q)) < 0.002] = 4                    # ambiguous bases
    w = rng.choice([5, 20, 100, 200], size=n).astype(np.int32)
    h0 = rng.integers(1, 200, size=n).astype(np.int32)
    return ql, qoff, q, tl, toff, t, w, h0


def _noalloc(lib, torch, p, opt, end_bonus, zdrop, max_qlen=None, stream=None):
    ql, qoff, q, tl, toff, t, w, h0 = p
    dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in p]
    out = torch.full((len(ql), 6), 12345, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    d = lambda x: C.c_void_p(x.data_ptr())
    rc = lib.b200_ksw_extend2_noalloc(
        0, C.c_int64(len(ql)), int(ql.max()) if max_qlen is None else max_qlen, d(dev[0]), d(dev[1]), d(dev[2]), d(dev[3]),
        d(dev[4]), d(dev[5]), opt.mat, opt.o_del, opt.e_del, opt.o_ins, opt.e_ins, d(dev[6]), end_bonus, zdrop, d(dev[7]),
        d(out), C.c_void_p(stream))
    torch.cuda.synchronize()
    return rc, out.cpu().numpy()


def test_ksw_extend2_noalloc_matches_batch_entry_point():
    torch = pytest.importorskip("torch")
    ctx = cm.context("cuda", "tiny")
    lib = ctx.lib
    lib.b200_noalloc_scratch_bytes.restype = C.c_int64
    assert lib.b200_noalloc_scratch_bytes() == 19398912
    assert lib.b200_noalloc_setup(0) == 0, lib.b200_last_error()
    opt = abi.default_opt()
    # call 1: many long-ish problems fill every warp's rows and run the work counter up
    p1 = _problems(101, 6000, 1022)
    rc, got = _noalloc(lib, torch, p1, opt, 5, 100)
    assert rc == 0, lib.b200_last_error()
    exp = ctx.ksw_extend2(*p1[:6], opt, p1[6], 5, 100, p1[7])
    assert np.array_equal(got, exp)
    # call 2, straight after: other problems, other z-drop / end bonus, a user stream, stale scratch underneath
    p2 = _problems(202, 3001, 300)
    s = torch.cuda.Stream()
    rc, got = _noalloc(lib, torch, p2, opt, 0, 0, stream=s.cuda_stream)
    assert rc == 0, lib.b200_last_error()
    exp = ctx.ksw_extend2(*p2[:6], opt, p2[6], 0, 0, p2[7])
    assert np.array_equal(got, exp)
    # golden tuples of the reference's ksw_extend2 (tests/golden/ksw_extend2.npz)
    import os
    g = np.load(os.path.join(cm.GOLDEN, "ksw_extend2.npz"))
    pg = tuple(g[k] for k in ("ql", "qoff", "q", "tl", "toff", "t", "w", "h0"))
    if int(pg[0].max()) <= 1022:
        rc, got = _noalloc(lib, torch, pg, opt, 5, 100)
        assert rc == 0 and np.array_equal(got, g["out_z100_eb5"])


def test_ksw_extend2_noalloc_rejects_sizes_beyond_the_scratch():
    torch = pytest.importorskip("torch")
    ctx = cm.context("cuda", "tiny")
    lib = ctx.lib
    opt = abi.default_opt()
    p = _problems(7, 64, 100)
    rc, _ = _noalloc(lib, torch, p, opt, 5, 100, max_qlen=1023)
    assert rc == -2 and b"1022" in lib.b200_last_error()
    # a problem that lies about its length is marked, not computed, and the others are unaffected
    ql = p[0].copy(); ql[3] = 5000
    h0 = p[7].copy(); h0[5] = 0
    rc, got = _noalloc(lib, torch, (ql,) + p[1:7] + (h0,), opt, 5, 100, max_qlen=100)
    assert rc == 0
    exp = ctx.ksw_extend2(*p[:6], opt, p[6], 5, 100, p[7])
    bad = np.zeros(64, dtype=bool); bad[[3, 5]] = True
    assert np.all(got[bad] == np.iinfo(np.int32).min) and np.array_equal(got[~bad], exp[~bad])
