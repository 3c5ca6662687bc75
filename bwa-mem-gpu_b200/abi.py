"""ctypes / numpy mirrors of the reference's ABI structs (this fork's layouts, /root/reference/bwa.h).

These are the types that cross the drop-in boundary:
  mem_opt_t     bwa.h:86-118   (168 B, `mat` at offset 136)
  mem_alnreg_t  bwa.h:145-163  (88 B)  -- the device->host payload of the hot path
  bwtintv_t     bwt.h:62-64    (32 B)
  mem_pestat_t  bwa.h:167-171
"""
from __future__ import annotations

import ctypes as C

import numpy as np


class MemOpt(C.Structure):
    _fields_ = [
        ("max_mem_intv", C.c_uint64),
        ("a", C.c_int), ("b", C.c_int),
        ("o_del", C.c_int), ("e_del", C.c_int),
        ("o_ins", C.c_int), ("e_ins", C.c_int),
        ("pen_unpaired", C.c_int),
        ("pen_clip5", C.c_int), ("pen_clip3", C.c_int),
        ("w", C.c_int),
        ("zdrop", C.c_int),
        ("T", C.c_int),
        ("flag", C.c_int),
        ("min_seed_len", C.c_int),
        ("min_chain_weight", C.c_int),
        ("max_chain_extend", C.c_int),
        ("split_factor", C.c_float),
        ("split_width", C.c_int),
        ("max_occ", C.c_int),
        ("max_chain_gap", C.c_int),
        ("n_threads", C.c_int),
        ("chunk_size", C.c_int),
        ("mask_level", C.c_float),
        ("drop_ratio", C.c_float),
        ("XA_drop_ratio", C.c_float),
        ("mask_level_redun", C.c_float),
        ("mapQ_coef_len", C.c_float),
        ("mapQ_coef_fac", C.c_int),
        ("max_ins", C.c_int),
        ("max_matesw", C.c_int),
        ("max_XA_hits", C.c_int), ("max_XA_hits_alt", C.c_int),
        ("mat", C.c_int8 * 25),
    ]


assert C.sizeof(MemOpt) == 168 and MemOpt.mat.offset == 136


class MemPestat(C.Structure):
    _fields_ = [("low", C.c_int), ("high", C.c_int), ("failed", C.c_int),
                ("avg", C.c_double), ("std", C.c_double)]


assert C.sizeof(MemPestat) == 32

# mem_alnreg_t as a numpy structured dtype (bit-fields n_comp:30,is_alt:2 share the last int32)
ALNREG_DTYPE = np.dtype([
    ("rb", "<i8"), ("re", "<i8"), ("hash", "<u8"), ("frac_rep", "<f4"),
    ("qb", "<i4"), ("qe", "<i4"), ("rid", "<i4"), ("score", "<i4"), ("truesc", "<i4"),
    ("sub", "<i4"), ("alt_sc", "<i4"), ("csub", "<i4"), ("sub_n", "<i4"), ("w", "<i4"),
    ("seedcov", "<i4"), ("secondary", "<i4"), ("secondary_all", "<i4"), ("seedlen0", "<i4"),
    ("n_comp_is_alt", "<u4"),
], align=True)
assert ALNREG_DTYPE.itemsize == 88

# mem_pestat_t (bwamem.h:59-63): int low, high, failed; double avg, std
PESTAT_DTYPE = np.dtype([("low", "<i4"), ("high", "<i4"), ("failed", "<i4"), ("avg", "<f8"), ("std", "<f8")], align=True)
assert PESTAT_DTYPE.itemsize == 32

# b200_aln_t (include/bwamem_b200.h): mem_aln_t without the floating-point fields
ALN_DTYPE = np.dtype([("pos", "<i8"), ("rid", "<i4"), ("flag", "<i4"), ("is_rev", "<i4"), ("is_alt", "<i4"),
                      ("n_cigar", "<i4"), ("NM", "<i4"), ("score", "<i4"), ("sub", "<i4"), ("alt_sc", "<i4"), ("l_md", "<i4"),
                      ("cigar_off", "<i8"), ("md_off", "<i8")], align=True)
assert ALN_DTYPE.itemsize == 64
SEL_DTYPE = np.dtype([("kind", "<i4"), ("mapq", "<i4"), ("xa_of", "<i4"), ("flag", "<i4")])      # b200_sel_t

INTV_DTYPE = np.dtype([("x0", "<u8"), ("x1", "<u8"), ("x2", "<u8"), ("info", "<u8")])

MEM_F_PE = 0x2


def default_opt() -> MemOpt:
    """mem_opt_init() defaults (bwamem.c:74-110) + bwa_fill_scmat (bwa.c:249)."""
    import math
    o = MemOpt()
    o.a, o.b = 1, 4
    o.o_del = o.o_ins = 6
    o.e_del = o.e_ins = 1
    o.w = 100
    o.T = 30
    o.zdrop = 100
    o.pen_unpaired = 17
    o.pen_clip5 = o.pen_clip3 = 5
    o.max_mem_intv = 20
    o.min_seed_len = 19
    o.split_width = 10
    o.max_occ = 500
    o.max_chain_gap = 10000
    o.max_ins = 10000
    o.mask_level = 0.50
    o.drop_ratio = 0.50
    o.XA_drop_ratio = 0.80
    o.split_factor = 1.5
    o.chunk_size = 30000000
    o.n_threads = 1
    o.max_XA_hits = 5
    o.max_XA_hits_alt = 200
    o.max_matesw = 50
    o.mask_level_redun = 0.95
    o.min_chain_weight = 0
    o.max_chain_extend = 1 << 30
    o.mapQ_coef_len = 50
    o.mapQ_coef_fac = int(math.log(50.0))
    fill_scmat(o)
    return o


def fill_scmat(o: MemOpt) -> None:
    k = 0
    for i in range(4):
        for j in range(4):
            o.mat[k] = o.a if i == j else -o.b
            k += 1
        o.mat[k] = -1
        k += 1
    for _ in range(5):
        o.mat[k] = -1
        k += 1
