"""ctypes binding of oracle/_ref/libbwaref.so (the unmodified reference CPU code + oracle/ref_shim.c).

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bwa_mem_gpu_b200.abi import ALNREG_DTYPE, INTV_DTYPE, MemOpt  # noqa: E402

REF_SO = os.path.join(ROOT, "oracle", "_ref", "libbwaref.so")


def available() -> bool:
    return os.path.exists(REF_SO)


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class RefLib:
    def __init__(self):
        self.lib = C.CDLL(REF_SO, mode=os.RTLD_LOCAL)
        L = self.lib
        L.ref_index_load.restype = C.c_void_p
        L.ref_index_load.argtypes = [C.c_char_p]
        L.ref_index_build.argtypes = [C.c_char_p, C.c_char_p]
        L.ref_index_destroy.argtypes = [C.c_void_p]
        L.ref_index_info.argtypes = [C.c_void_p, C.c_void_p]
        for f in ("ref_index_bwt", "ref_index_sa", "ref_index_pac", "ref_index_bwt_struct", "ref_index_bns_struct"):
            getattr(L, f).restype = C.c_void_p
            getattr(L, f).argtypes = [C.c_void_p]
        L.ref_index_ann_name.restype = C.c_char_p
        L.ref_index_ann_name.argtypes = [C.c_void_p, C.c_int]
        L.ref_opt_init.restype = C.c_void_p
        L.ref_collect_intv.restype = C.c_int64
        L.ref_align.restype = C.c_int64
        L.ref_process_seqs.restype = C.c_int64

    # ---------------- index
    def index_build(self, fa: str, prefix: str) -> None:
        rc = self.lib.ref_index_build(fa.encode(), prefix.encode())
        assert rc == 0

    def index_load(self, prefix: str) -> "RefIndex":
        h = self.lib.ref_index_load(prefix.encode())
        assert h, f"cannot load {prefix}"
        return RefIndex(self, h)

    def index_from_arrays(self, bwt_words, primary, L2, seq_len, sa, sa_intv, pac, l_pac, ann_off, ann_len) -> "RefIndex":
        """Wrap in-memory index arrays (kept alive by the returned object) into the reference's bwaidx_t."""
        keep = [np.ascontiguousarray(bwt_words, dtype=np.uint32), np.ascontiguousarray(sa, dtype=np.uint64),
                np.ascontiguousarray(pac, dtype=np.uint8), np.ascontiguousarray(ann_off, dtype=np.int64),
                np.ascontiguousarray(ann_len, dtype=np.int32), np.array(L2, dtype=np.uint64)]
        self.lib.ref_index_from_arrays.restype = C.c_void_p
        h = self.lib.ref_index_from_arrays(_p(keep[0], C.c_uint32), C.c_uint64(len(keep[0])), C.c_uint64(primary),
                                           _p(keep[5], C.c_uint64), C.c_uint64(seq_len), _p(keep[1], C.c_uint64),
                                           C.c_uint64(len(keep[1])), C.c_int(sa_intv), _p(keep[2], C.c_uint8),
                                           C.c_int64(l_pac), C.c_int(len(keep[3])), _p(keep[3], C.c_int64),
                                           _p(keep[4], C.c_int32))
        ri = RefIndex(self, h)
        ri._keep, ri._wrapped = keep, True
        return ri

    def opt_init(self) -> MemOpt:
        p = self.lib.ref_opt_init()
        o = MemOpt.from_buffer_copy(C.string_at(p, C.sizeof(MemOpt)))
        self.lib.ref_free(C.c_void_p(p))
        return o

    def sizeof(self, what: int) -> int:
        return self.lib.ref_sizeof(what)

    # ---------------- DP primitives
    def ksw_extend2(self, qlen, qoff, q, tlen, toff, t, opt: MemOpt, w, end_bonus, zdrop, h0):
        n = len(qlen)
        out = np.zeros((n, 6), dtype=np.int32)
        self.lib.ref_ksw_extend2(C.c_int64(n), _p(qlen, C.c_int32), _p(qoff, C.c_int64), _p(q, C.c_uint8),
                                 _p(tlen, C.c_int32), _p(toff, C.c_int64), _p(t, C.c_uint8),
                                 opt.mat, opt.o_del, opt.e_del, opt.o_ins, opt.e_ins,
                                 _p(w, C.c_int32), C.c_int(end_bonus), C.c_int(zdrop), _p(h0, C.c_int32),
                                 _p(out, C.c_int32))
        return out

    def ksw_align2(self, qlen, qoff, q, tlen, toff, t, opt: MemOpt, xtra):
        n = len(qlen)
        out = np.zeros((n, 7), dtype=np.int32)
        self.lib.ref_ksw_align2(C.c_int64(n), _p(qlen, C.c_int32), _p(qoff, C.c_int64), _p(q, C.c_uint8),
                                _p(tlen, C.c_int32), _p(toff, C.c_int64), _p(t, C.c_uint8),
                                opt.mat, opt.o_del, opt.e_del, opt.o_ins, opt.e_ins,
                                _p(xtra, C.c_int32), _p(out, C.c_int32))
        return out

    def ksw_global2_score(self, qlen, qoff, q, tlen, toff, t, opt: MemOpt, w):
        n = len(qlen)
        out = np.zeros(n, dtype=np.int32)
        self.lib.ref_ksw_global2(C.c_int64(n), _p(qlen, C.c_int32), _p(qoff, C.c_int64), _p(q, C.c_uint8),
                                 _p(tlen, C.c_int32), _p(toff, C.c_int64), _p(t, C.c_uint8),
                                 opt.mat, opt.o_del, opt.e_del, opt.o_ins, opt.e_ins,
                                 _p(w, C.c_int32), _p(out, C.c_int32), 0, None, None)
        return out


class RefIndex:
    def __init__(self, lib: RefLib, h):
        self.rl, self.lib, self.h = lib, lib.lib, C.c_void_p(h)
        info = np.zeros(16, dtype=np.int64)
        self.lib.ref_index_info(self.h, _p(info, C.c_int64))
        self.l_pac, self.seq_len, self.primary = int(info[0]), int(info[1]), int(info[2])
        self.L2 = [int(x) for x in info[3:8]]
        self.bwt_size, self.n_sa, self.sa_intv = int(info[8]), int(info[9]), int(info[10])
        self.n_seqs, self.n_holes = int(info[11]), int(info[12])

    def close(self):
        if self.h:
            if getattr(self, "_wrapped", False):
                self.lib.ref_index_release_wrapped(self.h)
            else:
                self.lib.ref_index_destroy(self.h)
            self.h = None

    # raw arrays (views over the reference's memory; copy if kept)
    def bwt_words(self):
        p = self.lib.ref_index_bwt(self.h)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), shape=(self.bwt_size,))

    def sa(self):
        p = self.lib.ref_index_sa(self.h)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint64)), shape=(self.n_sa,))

    def pac(self):
        p = self.lib.ref_index_pac(self.h)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(self.l_pac // 4 + 1,))

    def bwt_struct(self):
        return self.lib.ref_index_bwt_struct(self.h)

    def bns_struct(self):
        return self.lib.ref_index_bns_struct(self.h)

    def anns(self):
        off = np.zeros(self.n_seqs, dtype=np.int64)
        ln = np.zeros(self.n_seqs, dtype=np.int32)
        alt = np.zeros(self.n_seqs, dtype=np.int32)
        self.lib.ref_index_anns(self.h, _p(off, C.c_int64), _p(ln, C.c_int32), _p(alt, C.c_int32))
        names = [self.lib.ref_index_ann_name(self.h, i).decode() for i in range(self.n_seqs)]
        return off, ln, alt, names

    # FM-index primitives
    def occ4(self, k):
        k = np.ascontiguousarray(k, dtype=np.uint64)
        out = np.zeros((len(k), 4), dtype=np.uint64)
        self.lib.ref_occ4(self.h, C.c_int64(len(k)), _p(k, C.c_uint64), _p(out, C.c_uint64))
        return out

    def occ2x4(self, k, l):
        k = np.ascontiguousarray(k, dtype=np.uint64)
        l = np.ascontiguousarray(l, dtype=np.uint64)
        ok = np.zeros((len(k), 4), dtype=np.uint64)
        ol = np.zeros((len(k), 4), dtype=np.uint64)
        self.lib.ref_2occ4(self.h, C.c_int64(len(k)), _p(k, C.c_uint64), _p(l, C.c_uint64),
                           _p(ok, C.c_uint64), _p(ol, C.c_uint64))
        return ok, ol

    def extend(self, ik, is_back: int):
        ik = np.ascontiguousarray(ik, dtype=np.uint64)
        out = np.zeros((len(ik), 4, 3), dtype=np.uint64)
        self.lib.ref_extend(self.h, C.c_int64(len(ik)), _p(ik, C.c_uint64), C.c_int(is_back), _p(out, C.c_uint64))
        return out

    def sa_lookup(self, k):
        k = np.ascontiguousarray(k, dtype=np.uint64)
        out = np.zeros(len(k), dtype=np.uint64)
        self.lib.ref_sa(self.h, C.c_int64(len(k)), _p(k, C.c_uint64), _p(out, C.c_uint64))
        return out

    def smem1a(self, q, x, min_intv=1, max_intv=0, cap=4096):
        q = np.ascontiguousarray(q, dtype=np.uint8)
        out = np.zeros(cap, dtype=INTV_DTYPE)
        ret = C.c_int(0)
        n = self.lib.ref_smem1a(self.h, C.c_int(len(q)), _p(q, C.c_uint8), C.c_int(x), C.c_int(min_intv),
                                C.c_uint64(max_intv), C.c_int(cap), C.c_void_p(out.ctypes.data), C.byref(ret))
        assert n >= 0
        return out[:n].copy(), ret.value

    # batched stage seams
    def collect_intv(self, opt: MemOpt, seqs, off, cap_per_read=256):
        n = len(off) - 1
        cap = n * cap_per_read
        out = np.zeros(cap, dtype=INTV_DTYPE)
        cnt = np.zeros(n, dtype=np.int32)
        tot = self.lib.ref_collect_intv(self.h, C.byref(opt), C.c_int(n), _p(seqs, C.c_uint8), _p(off, C.c_int64),
                                        C.c_int64(cap), C.c_void_p(out.ctypes.data), _p(cnt, C.c_int32))
        assert tot >= 0
        return out[:tot].copy(), cnt

    def chains(self, opt: MemOpt, seqs, off, stage=1, cap_chn_per_read=64, cap_seed_per_read=1024):
        n = len(off) - 1
        cc, cs = n * cap_chn_per_read, n * cap_seed_per_read
        chn = np.zeros((cc, 8), dtype=np.int64)
        sd = np.zeros((cs, 4), dtype=np.int64)
        nch = np.zeros(n, dtype=np.int32)
        tot = np.zeros(2, dtype=np.int64)
        rc = self.lib.ref_chains(self.h, C.byref(opt), C.c_int(stage), C.c_int(n), _p(seqs, C.c_uint8),
                                 _p(off, C.c_int64), C.c_int64(cc), _p(chn, C.c_int64), _p(nch, C.c_int32),
                                 C.c_int64(cs), _p(sd, C.c_int64), _p(tot, C.c_int64))
        assert rc == 0, "chain buffers too small"
        return chn[:tot[0]].copy(), nch, sd[:tot[1]].copy()

    def align(self, opt: MemOpt, seqs, off, n_threads=1, cap_per_read=64, count_only=False):
        n = len(off) - 1
        cnt = np.zeros(n, dtype=np.int32)
        if count_only:
            tot = self.lib.ref_align(self.h, C.byref(opt), C.c_int(n), _p(seqs, C.c_uint8), _p(off, C.c_int64),
                                     C.c_int(n_threads), C.c_int64(0), None, _p(cnt, C.c_int32))
            return None, cnt
        cap = n * cap_per_read
        out = np.zeros(cap, dtype=ALNREG_DTYPE)
        tot = self.lib.ref_align(self.h, C.byref(opt), C.c_int(n), _p(seqs, C.c_uint8), _p(off, C.c_int64),
                                 C.c_int(n_threads), C.c_int64(cap), C.c_void_p(out.ctypes.data), _p(cnt, C.c_int32))
        assert tot >= 0
        return out[:tot].copy(), cnt

    def reg2aln(self, opt: MemOpt, seqs, off, regs, cnt):
        """mem_reg2aln (bwamem.c:1099) for every region -> (n x 8 int64: pos rid is_rev n_cigar NM score sub l_md, cigar, md)."""
        n = len(off) - 1
        regs = np.ascontiguousarray(regs); cnt = np.ascontiguousarray(cnt, dtype=np.int32)
        aln = np.zeros((max(1, len(regs)), 8), dtype=np.int64)
        tot = (C.c_int64 * 2)()
        cap_c, cap_m = 8 * len(regs) + 64, 24 * len(regs) + 64
        self.lib.ref_reg2aln.restype = C.c_int64
        while True:
            cig = np.zeros(cap_c, dtype=np.uint32); md = np.zeros(cap_m, dtype=np.uint8)
            rc = self.lib.ref_reg2aln(self.h, C.byref(opt), C.c_int(n), _p(seqs, C.c_uint8), _p(off, C.c_int64),
                                      C.c_void_p(regs.ctypes.data), _p(cnt, C.c_int32), C.c_void_p(aln.ctypes.data),
                                      C.c_int64(cap_c), _p(cig, C.c_uint32), C.c_int64(cap_m), C.c_void_p(md.ctypes.data), tot)
            if rc < 0:
                cap_c, cap_m = int(tot[0]) + 64, int(tot[1]) + 64
                continue
            return aln[:rc], cig[: int(tot[0])], md[: int(tot[1])]

    def reg2sam_se(self, opt: MemOpt, seqs_nt4, off, regs, cnt, n_processed=0):
        """worker2, SE (bwamem.c:1195-1205): mem_mark_primary_se (+ mem_reorder_primary5), mem_approx_mapq_se per region,
        mem_reg2sam -> (regions as updated, mapq per region, SAM text bytes); reads are named r<global index>."""
        n = len(off) - 1
        regs = np.array(regs, copy=True); cnt = np.ascontiguousarray(cnt, dtype=np.int32)
        mapq = np.zeros(max(1, len(regs)), dtype=np.int32)
        cap = 600 * max(1, n) + 4 * int(off[-1]) + 400 * len(regs) + 1024
        self.lib.ref_reg2sam_se.restype = C.c_int64
        while True:
            sam = np.zeros(cap, dtype=np.uint8)
            work = regs.copy()
            need = self.lib.ref_reg2sam_se(self.h, C.byref(opt), C.c_int(n), C.c_int64(n_processed), _p(seqs_nt4, C.c_uint8),
                                           _p(off, C.c_int64), C.c_void_p(work.ctypes.data), _p(cnt, C.c_int32),
                                           _p(mapq, C.c_int32), C.c_int64(cap), C.c_void_p(sam.ctypes.data))
            if need > cap:
                cap = int(need) + 1024
                continue
            return work, mapq[: len(regs)], sam[:need].tobytes()

    def pestat(self, opt: MemOpt, regs, cnt):
        """mem_pestat (bwamem_pair.c:72) -> 4 x PESTAT_DTYPE"""
        from bwa_mem_gpu_b200.abi import PESTAT_DTYPE
        pes = np.zeros(4, dtype=PESTAT_DTYPE)
        regs = np.ascontiguousarray(regs); cnt = np.ascontiguousarray(cnt, dtype=np.int32)
        self.lib.ref_pestat(self.h, C.byref(opt), C.c_int(len(cnt)), C.c_void_p(regs.ctypes.data), _p(cnt, C.c_int32),
                            C.c_void_p(pes.ctypes.data))
        return pes

    def mate_rescue(self, opt: MemOpt, pes, seqs, off, regs, cnt):
        """The rescue block of mem_sam_pe (bwamem_pair.c:291-302) for every pair -> (regs, counts, n_sw)."""
        n = len(off) - 1
        regs = np.ascontiguousarray(regs); cnt = np.ascontiguousarray(cnt, dtype=np.int32)
        pes = np.ascontiguousarray(pes)
        cap = len(regs) + 8 * n + 64
        out = np.zeros(cap, dtype=ALNREG_DTYPE)
        ocnt = np.zeros(max(n, 1), dtype=np.int32)
        nsw = C.c_int64(0)
        self.lib.ref_mate_rescue.restype = C.c_int64
        tot = self.lib.ref_mate_rescue(self.h, C.byref(opt), C.c_void_p(pes.ctypes.data), C.c_int(n), _p(seqs, C.c_uint8),
                                       _p(off, C.c_int64), C.c_void_p(regs.ctypes.data), _p(cnt, C.c_int32), C.c_int64(cap),
                                       C.c_void_p(out.ctypes.data), _p(ocnt, C.c_int32), C.byref(nsw))
        assert tot >= 0
        return out[:tot].copy(), ocnt[:n], nsw.value

    def process_seqs(self, opt: MemOpt, seqs, off, names=None, quals=None, n_processed=0, pes0=None, comments=None, rg=None):
        n = len(off) - 1
        if names is None:
            names = [f"r{n_processed + i}" for i in range(n)]
        nm = b"\0".join(s.encode() for s in names) + b"\0"
        cm = b"\0".join((c or "").encode() for c in comments) + b"\0" if comments is not None else None
        cap = int(off[-1]) * 6 + n * 512 + 4096
        buf = C.create_string_buffer(cap)
        self.lib.ref_process_seqs2.restype = C.c_int64
        tot = self.lib.ref_process_seqs2(self.h, C.byref(opt), C.c_int64(n_processed), C.c_int(n), nm, cm,
                                         _p(seqs, C.c_uint8), _p(off, C.c_int64),
                                         _p(quals, C.c_uint8) if quals is not None else None, rg.encode() if rg else None,
                                         (C.c_void_p(pes0.ctypes.data) if isinstance(pes0, np.ndarray) else C.byref(pes0)) if pes0 is not None else None,
                                         C.c_int64(cap), buf)
        assert tot >= 0, f"SAM buffer too small ({-tot})"
        self.lib.ref_last_process_seqs_seconds.restype = C.c_double
        self.last_process_seqs_s = self.lib.ref_last_process_seqs_seconds()      # the mem_process_seqs call alone
        return buf.raw[:tot]
