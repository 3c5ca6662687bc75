"""Pins the plain-C restatement under oracle/ (ora_fmindex.c, ora_ksw.c -> oracle/_build/liboracle.so) against the
committed golden vectors (outputs of the unmodified reference) and, where oracle/_ref is present, against the
reference itself on further inputs.  The index arrays come from our own host-side builder (the same arrays the
reference's `bwa index` writes, tests/test_index.py), so this file needs neither a GPU nor /root/reference."""
import ctypes as C
import os

import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import abi, lib as L

ORACLE_SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_build", "liboracle.so")
pytestmark = pytest.mark.skipif(not os.path.exists(ORACLE_SO), reason="oracle/_build/liboracle.so not built (make -C oracle oracle)")
G = "tiny"


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class Ora:
    def __init__(self, genome, host_lib=None):
        """genome: a name known to tests/common.py or a list of contigs (nt4 arrays)"""
        self.lib = C.CDLL(ORACLE_SO)
        contigs = cm.genome(genome) if isinstance(genome, str) else genome
        self.h = L.HostIndex(contigs, host_lib or cm.backend_lib("emu"))      # host-side index builder only
        self.bwt, self.sa = np.ascontiguousarray(self.h.bwt_words()), np.ascontiguousarray(self.h.sa())
        l2 = (C.c_uint64 * 5)(*[int(x) for x in self.h.L2])
        self.lib.ora_index_new.restype = C.c_void_p
        self.ix = C.c_void_p(self.lib.ora_index_new(_p(self.bwt, C.c_uint32), C.c_uint64(self.h.primary), l2,
                                                    C.c_uint64(self.h.seq_len), _p(self.sa, C.c_uint64),
                                                    C.c_uint64(len(self.sa)), C.c_int(self.h.sa_intv)))
        self.lib.ora_sa.restype = C.c_uint64

    def occ4(self, k):
        out = np.zeros((len(k), 4), np.uint64)
        for i, v in enumerate(k):
            self.lib.ora_occ4(self.ix, C.c_uint64(int(v)), _p(out[i], C.c_uint64))
        return out

    def extend(self, ik, back):
        out = np.zeros((len(ik), 4, 3), np.uint64)
        kids = (C.c_uint64 * 16)()
        for i in range(len(ik)):
            iv = (C.c_uint64 * 4)(int(ik[i, 0]), int(ik[i, 1]), int(ik[i, 2]), 0)
            self.lib.ora_extend(self.ix, iv, kids, C.c_int(back))
            out[i] = np.array(list(kids), np.uint64).reshape(4, 4)[:, :3]
        return out

    def smem1a(self, q, x, min_intv, max_intv):
        q = np.ascontiguousarray(q, np.uint8)
        out = np.zeros((len(q) + 2, 4), np.uint64)
        nxt = C.c_int(0)
        n = self.lib.ora_smem1a(self.ix, C.c_int(len(q)), _p(q, C.c_uint8), C.c_int(x), C.c_int(min_intv), C.c_uint64(max_intv),
                                C.c_int(len(q) + 2), C.c_void_p(out.ctypes.data), C.byref(nxt))
        assert n >= 0
        return out[:n].copy(), nxt.value

    def collect_intv(self, opt, seqs, off):
        ivs, cnt = [], []
        for r in range(len(off) - 1):
            q = np.ascontiguousarray(seqs[off[r]:off[r + 1]], np.uint8)
            out = np.zeros((4 * len(q) + 16, 4), np.uint64)
            n = 0
            if len(q) >= opt.min_seed_len:                                    # mem_chain's length filter, bwamem.c:267
                n = self.lib.ora_collect_intv(self.ix, C.c_int(len(q)), _p(q, C.c_uint8), C.c_int(opt.min_seed_len),
                                              C.c_float(opt.split_factor), C.c_int(opt.split_width), C.c_int(int(opt.max_mem_intv)),
                                              C.c_int(len(out)), C.c_void_p(out.ctypes.data))
                assert n >= 0
            ivs.append(out[:n]); cnt.append(n)
        return np.concatenate(ivs), np.array(cnt, np.int32)

    def ksw_extend2(self, ql, qoff, q, tl, toff, t, opt, w, end_bonus, zdrop, h0):
        out = np.zeros((len(ql), 6), np.int32)
        self.lib.ora_ksw_extend2_batch(C.c_int64(len(ql)), _p(ql, C.c_int32), _p(qoff, C.c_int64), _p(q, C.c_uint8), _p(tl, C.c_int32),
                                       _p(toff, C.c_int64), _p(t, C.c_uint8), opt.mat, opt.o_del, opt.e_del, opt.o_ins, opt.e_ins,
                                       _p(np.ascontiguousarray(w, np.int32), C.c_int32), C.c_int(end_bonus), C.c_int(zdrop),
                                       _p(np.ascontiguousarray(h0, np.int32), C.c_int32), _p(out, C.c_int32))
        return out


@pytest.fixture(scope="module")
def ora():
    return Ora(G)


def _golden(name):
    return np.load(cm.golden_path(name))


def test_oracle_fmindex_golden(ora):
    g = _golden("fmindex.npz")
    assert np.array_equal(ora.occ4(g["k"]), g["occ4"])
    assert np.array_equal(ora.occ4(g["l"]), g["occl"]) and np.array_equal(ora.occ4(g["k"]), g["occk"])
    assert np.array_equal(ora.extend(g["ik"], 0), g["ext_fwd"])
    assert np.array_equal(ora.extend(g["ik"], 1), g["ext_back"])
    assert np.array_equal(ora.extend(g["ik2"], 1), g["ext2_back"])
    sa = np.array([ora.lib.ora_sa(ora.ix, C.c_uint64(int(k))) for k in g["sa_k"]], np.uint64)
    assert np.array_equal(sa, g["sa"])


def test_oracle_smem_golden(ora):
    g = _golden("smem1a.npz")
    seqs, off = cm.reads_for(G, 300, seed=11, n_rate=0.004)
    for max_intv in (0, 20):
        ivs, cnts, rets = [], [], []
        for r in range(300):
            iv, ret = ora.smem1a(seqs[off[r]:off[r + 1]], int(g["xs"][r]), int(g["mi"][r]), max_intv)
            ivs.append(iv); cnts.append(len(iv)); rets.append(ret)
        assert np.array_equal(np.array(cnts, np.int32), g[f"cnt{max_intv}"])
        assert np.array_equal(np.array(rets, np.int32), g[f"ret{max_intv}"])
        assert np.array_equal(np.concatenate(ivs), g[f"iv{max_intv}"])
    g = _golden("collect_intv.npz")
    seqs, off = cm.reads_for(G, 600, seed=21, n_rate=0.001)
    iv, cnt = ora.collect_intv(abi.default_opt(), seqs, off)
    assert np.array_equal(cnt, g["cnt"]) and np.array_equal(iv, g["iv"])


def test_oracle_ksw_extend2_golden(ora):
    g = _golden("ksw_extend2.npz")
    args = [g[k] for k in ("ql", "qoff", "q", "tl", "toff", "t")]
    opt = abi.default_opt()
    assert np.array_equal(ora.ksw_extend2(*args, opt, g["w"], 5, 100, g["h0"]), g["out_z100_eb5"])
    assert np.array_equal(ora.ksw_extend2(*args, opt, g["w"], 0, 0, g["h0"]), g["out_z0_eb0"])


@pytest.mark.skipif(not cm.have_ref(), reason="oracle/_ref not built")
def test_oracle_vs_reference_more_inputs():
    """Beyond the golden vectors: repeat-rich genome, ambiguous bases, non-default scoring."""
    from test_stages import _ksw_problems
    o = Ora("rep")
    idx = cm.ref_index("rep")
    opt = abi.default_opt()
    seqs, off = cm.reads_for("rep", 200, seed=5, n_rate=0.01, indel=0.01)
    a = idx.collect_intv(opt, seqs, off)
    b = o.collect_intv(opt, seqs, off)
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[0].view(np.uint64).reshape(-1, 4), b[0])
    rl = cm.reflib()
    rng = np.random.default_rng(3)
    for (zdrop, eb, a_, b_, od, oi, ed, ei) in [(100, 5, 1, 4, 6, 6, 1, 1), (10, 0, 2, 3, 4, 5, 2, 1), (0, 7, 1, 1, 1, 2, 1, 1)]:
        opt = abi.default_opt()
        opt.a, opt.b, opt.o_del, opt.o_ins, opt.e_del, opt.e_ins = a_, b_, od, oi, ed, ei
        abi.fill_scmat(opt)
        ql, qoff, q, tl, toff, t = _ksw_problems(rng, 400)
        w = rng.choice([1, 5, 30, 100, 200], size=400).astype(np.int32)
        h0 = rng.integers(1, 200, size=400).astype(np.int32)
        assert np.array_equal(rl.ksw_extend2(ql, qoff, q, tl, toff, t, opt, w, eb, zdrop, h0),
                              o.ksw_extend2(ql, qoff, q, tl, toff, t, opt, w, eb, zdrop, h0))
