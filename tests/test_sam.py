"""End-to-end drop-in parity: our mem_process_seqs (worker1, mem_pestat, mate rescue, worker2 and the SAM text on the device;
optionally the reference's own worker2 host code) must produce byte-identical SAM records to the reference's mem_process_seqs (bwamem.c:1215) --
POS, CIGAR, MAPQ, NM, MD, AS, XS, SA/XA tags, flags -- for single-end and paired-end batches.
"""
import ctypes as C
import os

import numpy as np
import pytest

import common as cm
import refapi
from bwa_mem_gpu_b200 import abi, synth, lib as L

BACKENDS = [pytest.param("emu", id="emu"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]
needs_ref = pytest.mark.skipif(not cm.have_ref(), reason="oracle/_ref not built")


class BSeq1(C.Structure):          # bseq1_t, bwa.h:56-63
    _fields_ = [("l_seq", C.c_int), ("id", C.c_int), ("name", C.c_char_p), ("comment", C.c_char_p),
                ("seq", C.POINTER(C.c_char)), ("qual", C.c_char_p), ("sam", C.c_void_p),
                ("l_name", C.c_int8), ("l_comment", C.c_int8), ("l_qual", C.c_int16)]


def _our_sam(backend, gname, opt, seqs, off, n_processed=0, names=None, n_shards=1, quals=None, host_worker2=-1):
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    C.CDLL(refapi.REF_SO, mode=os.RTLD_GLOBAL)          # the "host program": the reference's worker2 functions
    lib = ctx.lib
    n = len(off) - 1
    arr = (BSeq1 * n)()
    keep = []
    asc = synth.to_ascii(seqs)
    for i in range(n):
        s = C.create_string_buffer(bytes(asc[off[i]:off[i + 1]]) + b"\0")
        nm = (names[i] if names else f"r{n_processed + i}").encode()
        keep.append((s, nm))
        arr[i].l_seq = int(off[i + 1] - off[i]); arr[i].id = i
        arr[i].name = nm; arr[i].comment = None; arr[i].seq = C.cast(s, C.POINTER(C.c_char)); arr[i].qual = None
        if quals is not None:
            q = bytes(quals[off[i]:off[i + 1]]) + b"\0"
            keep.append(q)
            arr[i].qual = q
    if n_shards == 1:
        lib.b200_process_seqs_set_ctx(ctx.h)
    else:                                                   # several contexts (here: all on device 0), one block of reads each
        extra = []
        for k in range(n_shards - 1):
            try:                                            # a second physical GPU when there is one
                c = L.Context((k + 1) % 2 if backend == "cuda" else 0, lib)
            except L.B200Error:
                c = L.Context(0, lib)
            c.upload_host_index(ctx._host_index)
            extra.append(c)
        keep.append(extra)
        hs = (C.c_void_p * n_shards)(ctx.h, *[c.h for c in extra])
        lib.b200_process_seqs_set_ctxs(n_shards, hs)
    lib.b200_process_seqs_host_worker2(host_worker2)
    lib.b200_mem_process_seqs(C.byref(opt), C.c_void_p(idx.bwt_struct()), C.c_void_p(idx.bns_struct()),
                              C.c_void_p(idx.pac().ctypes.data), C.c_int64(n_processed), C.c_int(n), arr, None)
    out = b"".join(C.string_at(arr[i].sam) for i in range(n))
    for i in range(n):
        idx.lib.ref_free(C.c_void_p(arr[i].sam))
    return out


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname", ["small", "rep", "dup"])
def test_sam_single_end(backend, gname):
    """Reads -> SAM text through the reference's batch entry.  worker2 on the device (the default) and through the host
    program's functions; "dup": duplicated segments and an ALT contig (XA, pa, secondary_all); with qualities."""
    idx = cm.ref_index(gname)
    opt = abi.default_opt()
    opt.n_threads = 4
    n = 400 if backend == "emu" else 20000
    seqs, off = cm.reads_for(gname, n, seed=71, indel=0.004, sub=0.015)
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, n_processed=1000)
    got = _our_sam(backend, gname, opt, seqs, off, n_processed=1000, host_worker2=-1)      # auto: short reads -> device worker2
    assert got == exp
    assert exp.count(b"\n") >= n
    assert _our_sam(backend, gname, opt, seqs, off, n_processed=1000, host_worker2=0) == exp
    assert _our_sam(backend, gname, opt, seqs, off, n_processed=1000, host_worker2=1) == exp
    quals = np.random.default_rng(5).integers(35, 74, size=len(seqs), dtype=np.uint8)
    seqs2, off2 = cm.ragged_reads(gname)
    o = abi.default_opt()
    o.n_threads = 2
    o.flag |= 0x800 | 0x200                                 # -5 -Y
    for oo in (opt, o):
        exp = idx.process_seqs(oo, synth.to_ascii(seqs2), off2, quals=quals[: len(seqs2)], n_processed=5)
        assert _our_sam(backend, gname, oo, seqs2, off2, n_processed=5, quals=quals[: len(seqs2)]) == exp


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("damage", ["window", "periodic"])
def test_sam_paired_end(backend, damage):
    """Interleaved pairs (MEM_F_PE) through the drop-in: worker1, mate rescue, mem_pair / MAPQ / flags and the paired SAM text
    on the device, the batch's mem_pestat as a host reduction; then the older route with the host program's mem_pestat and
    mem_sam_pe; "periodic" damage leaves mates that only mem_matesw can place."""
    gname = "small"
    idx = cm.ref_index(gname)
    opt = abi.default_opt()
    opt.flag |= 0x2
    opt.n_threads = 4
    npairs = 300 if backend == "emu" else 8000
    reads, _ = synth.simulate_pairs(cm.genome(gname), npairs, seed=81, damaged_frac=0.1, damage=damage)
    seqs = np.ascontiguousarray(reads.reshape(-1))
    off = np.arange(0, 2 * npairs + 1, dtype=np.int64) * reads.shape[1]
    names = [f"p{i // 2}" for i in range(2 * npairs)]
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, names=names, n_processed=0)
    got = _our_sam(backend, gname, opt, seqs, off, n_processed=0, names=names)                  # everything on the device
    assert got == exp
    assert exp.count(b"\n") >= 2 * npairs
    assert _our_sam(backend, gname, opt, seqs, off, n_processed=0, names=names, host_worker2=1) == exp      # the host program's mem_sam_pe
    opt.flag |= 0x800 | 0x200                               # -5 -Y on pairs (Hi-C usage): mem_sam_pe reorders itself (bwamem_pair.c:305-307)
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, names=names, n_processed=0)
    assert _our_sam(backend, gname, opt, seqs, off, n_processed=0, names=names) == exp


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("pe", [0, 1])
def test_sam_sharded_over_contexts(backend, pe):
    """One batch cut into 3 contiguous blocks (whole pairs) on 3 contexts: SAM identical to the reference's, i.e. to
    the unsharded run; the PE case needs the batch-level mem_pestat over all blocks before each block's rescue."""
    gname = "small"
    idx = cm.ref_index(gname)
    opt = abi.default_opt()
    opt.n_threads = 4
    if pe:
        opt.flag |= 0x2
        npairs = 301 if backend == "emu" else 6001
        reads, _ = synth.simulate_pairs(cm.genome(gname), npairs, seed=83, damaged_frac=0.1, damage="periodic")
        seqs = np.ascontiguousarray(reads.reshape(-1))
        off = np.arange(0, 2 * npairs + 1, dtype=np.int64) * reads.shape[1]
        names = [f"p{i // 2}" for i in range(2 * npairs)]
    else:
        seqs, off = cm.reads_for(gname, 401 if backend == "emu" else 20001, seed=73, indel=0.004, sub=0.015)
        names = None
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, names=names, n_processed=64)
    got = _our_sam(backend, gname, opt, seqs, off, n_processed=64, names=names, n_shards=3)
    assert got == exp
