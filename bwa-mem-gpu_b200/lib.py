"""ctypes binding of the C ABI in include/bwamem_b200.h.

The product library is `libbwamem_b200.so` (nvcc, sm_100a), in-tree next to this file.  There is NO
CPU implementation behind this binding: `load()` raises if the CUDA library is missing, and
`Context()` raises if no CUDA device is present.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .abi import ALN_DTYPE, ALNREG_DTYPE, INTV_DTYPE, MemOpt

HERE = os.path.dirname(os.path.abspath(__file__))
PRODUCT_SO = os.path.join(HERE, "libbwamem_b200.so")

IDX_FULL_SA = 0x1
E_FAIL, E_CAPACITY = -(1 << 63) + 1, -(1 << 63) + 2      # B200_E_FAIL / B200_E_CAPACITY (include/bwamem_b200.h)

_SYMBOLS = [
    "b200_ctx_create", "b200_ctx_destroy", "b200_ctx_tune", "b200_ctx_trim", "b200_microbench", "b200_last_error", "b200_index_upload", "b200_index_upload_raw",
    "b200_occ4_batch", "b200_2occ4_batch", "b200_extend_batch", "b200_sa_batch", "b200_smem1a_batch",
    "b200_collect_intv_batch", "b200_chain_batch", "b200_ksw_extend2_batch", "b200_ksw_align2_batch", "b200_align_batch",
    "b200_batch_upload", "b200_batch_run", "b200_batch_download", "b200_batch_counters", "b200_batch_rescue", "b200_batch_set_regs",
    "b200_hostidx_build", "b200_hostidx_free", "b200_hostidx_info", "b200_hostidx_bwt", "b200_hostidx_sa",
    "b200_hostidx_pac", "b200_index_upload_host", "b200_index_build_device", "b200_index_info", "b200_index_download", "b200_reg2aln_batch", "b200_reg2sam_se_batch", "b200_sam_se_format", "b200_align_sam_se_batch", "b200_host_alloc", "b200_host_free", "b200_batch_reg2sam_se", "b200_process_seqs_host_worker2", "b200_process_seqs_set_rg_id",
    "b200_reads_to_sam", "b200_index_replicate", "b200_index_save", "b200_index_load", "b200_fastq_open", "b200_fastq_read", "b200_fastq_close", "b200_fastq_to_sam", "b200_ksw_extend2_noalloc", "b200_noalloc_scratch_bytes", "b200_noalloc_setup",
    "b200_process_seqs_set_ctx", "b200_process_seqs_set_ctxs", "b200_process_seqs_use_devices", "b200_mem_process_seqs", "mem_process_seqs",
    "bwt_occ4", "bwt_2occ4", "bwt_extend", "bwt_sa", "bwt_smem1a", "ksw_extend2", "mem_align1_core",
]


def exported_symbols():
    return list(_SYMBOLS)


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def load(path: str | None = None) -> C.CDLL:
    path = path or PRODUCT_SO
    if not os.path.exists(path):
        raise RuntimeError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
    lib = C.CDLL(path, mode=os.RTLD_LOCAL)
    lib.b200_ctx_create.restype = C.c_void_p
    lib.b200_ctx_create.argtypes = [C.c_int]
    lib.b200_ctx_destroy.argtypes = [C.c_void_p]
    lib.b200_last_error.restype = C.c_char_p
    lib.b200_collect_intv_batch.restype = C.c_int64
    lib.b200_align_batch.restype = C.c_int64
    lib.b200_batch_download.restype = C.c_int64
    lib.b200_reg2aln_batch.restype = C.c_int64
    lib.b200_reg2sam_se_batch.restype = C.c_int64
    lib.b200_sam_se_format.restype = C.c_int64
    lib.b200_align_sam_se_batch.restype = C.c_int64
    lib.b200_reads_to_sam.restype = C.c_int64
    lib.b200_host_alloc.restype = C.c_void_p
    lib.b200_host_alloc.argtypes = [C.c_size_t]
    lib.b200_host_free.argtypes = [C.c_void_p]
    lib.b200_batch_reg2sam_se.restype = C.c_int64
    lib.b200_hostidx_build.restype = C.c_void_p
    lib.b200_hostidx_build.argtypes = [C.c_void_p, C.c_int64]
    lib.b200_hostidx_free.argtypes = [C.c_void_p]
    lib.b200_hostidx_info.argtypes = [C.c_void_p, C.c_void_p]
    for f in ("b200_hostidx_bwt", "b200_hostidx_sa", "b200_hostidx_pac"):
        getattr(lib, f).restype = C.c_void_p
        getattr(lib, f).argtypes = [C.c_void_p]
    return lib


class HostIndex:
    """In-memory index built by our own SA-IS builder (the arrays `bwa index` would write)."""

    def __init__(self, contigs, lib: C.CDLL | None = None, names=None):
        self.lib = lib or load()
        fwd = np.ascontiguousarray(np.concatenate(contigs), dtype=np.uint8)
        self.l_pac = int(len(fwd))
        h = self.lib.b200_hostidx_build(C.c_void_p(fwd.ctypes.data), C.c_int64(self.l_pac))
        if not h:
            raise B200Error("index build failed (bases must be nt4 codes 0..3)")
        self.h = C.c_void_p(h)
        info = np.zeros(11, dtype=np.uint64)
        self.lib.b200_hostidx_info(self.h, C.c_void_p(info.ctypes.data))
        self.primary = int(info[0])
        self.L2 = [int(x) for x in info[1:6]]
        self.seq_len, self.bwt_size, self.n_sa, self.sa_intv = int(info[6]), int(info[7]), int(info[8]), int(info[9])
        lens = np.array([len(c) for c in contigs], dtype=np.int64)
        self.ann_off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
        self.ann_len = lens.astype(np.int32)
        self.ann_alt = np.zeros(len(contigs), dtype=np.int32)
        self.names = names or [f"chr{i + 1}" for i in range(len(contigs))]

    def bwt_words(self):
        p = self.lib.b200_hostidx_bwt(self.h)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), shape=(self.bwt_size,))

    def sa(self):
        p = self.lib.b200_hostidx_sa(self.h)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint64)), shape=(self.n_sa,))

    def pac(self):
        p = self.lib.b200_hostidx_pac(self.h)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(self.l_pac // 4 + 1,))

    def close(self):
        if self.h:
            self.lib.b200_hostidx_free(self.h)
            self.h = None


class B200Error(RuntimeError):
    pass


class ReadsT(C.Structure):          # b200_reads_t
    _fields_ = [("n_reads", C.c_int64), ("seq", C.c_void_p), ("seq_off", C.c_void_p), ("qual", C.c_void_p),
                ("name", C.c_void_p), ("name_off", C.c_void_p), ("comment", C.c_void_p), ("comment_off", C.c_void_p)]


class FastqReader:
    """b200_fastq_*: batches of reads from FASTQ / FASTA files (plain, gzip, BGZF) as numpy views of the reader's buffers."""

    def __init__(self, fn1, fn2=None, n_threads=8, lib=None):
        self.lib = lib or load()
        self.lib.b200_fastq_open.restype = C.c_void_p
        self.lib.b200_fastq_read.restype = C.c_int64
        h = self.lib.b200_fastq_open(fn1.encode(), fn2.encode() if fn2 else None, C.c_int(n_threads))
        if not h:
            raise B200Error(self.lib.b200_last_error().decode())
        self.h = C.c_void_p(h)

    def read(self, chunk_bases):
        """Next batch: dict with n, seqs, off, quals, names (flat, off), comments ((flat, off) or None); None at the end."""
        R = ReadsT()
        n = self.lib.b200_fastq_read(self.h, C.c_int64(chunk_bases), C.byref(R))
        if n < 0:
            raise B200Error(self.lib.b200_last_error().decode())
        if n == 0:
            return None
        view = lambda p, cnt, ty: np.ctypeslib.as_array(C.cast(C.c_void_p(p), C.POINTER(ty)), shape=(max(1, cnt),))[:cnt]
        off = view(R.seq_off, n + 1, C.c_int64)
        noff = view(R.name_off, n + 1, C.c_int64)
        out = {"n": int(n), "off": off, "seqs": view(R.seq, int(off[n]), C.c_uint8),
               "quals": view(R.qual, int(off[n]), C.c_uint8) if R.qual else None,
               "names": (view(R.name, int(noff[n]), C.c_uint8), noff), "comments": None, "struct": R}
        if R.comment:
            coff = view(R.comment_off, n + 1, C.c_int64)
            out["comments"] = (view(R.comment, int(coff[n]), C.c_uint8), coff)
        return out

    def close(self):
        if self.h:
            self.lib.b200_fastq_close(self.h)
            self.h = None


def fastq_to_sam(ctxs, opt, ann_names, fn1, fn2=None, chunk_bases=100_000_000, out_path=None, n_threads=8, copy_comment=False, rg=None):
    """b200_fastq_to_sam: FASTQ file(s) -> SAM records appended to out_path (None: discarded).  Returns (reads, stats)."""
    lib = ctxs[0].lib
    lib.b200_fastq_to_sam.restype = C.c_int64
    a_names = (C.c_char_p * max(1, len(ann_names)))(*[s.encode() for s in ann_names])
    hs = (C.c_void_p * len(ctxs))(*[c.h for c in ctxs])
    stats = (C.c_double * 8)()
    fd = os.open(out_path, os.O_WRONLY | os.O_CREAT | os.O_APPEND, 0o644) if out_path else -1
    try:
        n = lib.b200_fastq_to_sam(hs, len(ctxs), C.byref(opt), a_names, None, fn1.encode(), fn2.encode() if fn2 else None,
                                  C.c_int64(chunk_bases), None, rg.encode() if rg else None, C.c_int(n_threads),
                                  C.c_int(1 if copy_comment else 0), C.c_int(fd), stats)
    finally:
        if fd >= 0:
            os.close(fd)
    ctxs[0]._chk(n, "fastq_to_sam")
    return n, list(stats)


def flat_strings(strs):
    """list of str / bytes -> (uint8 array, int64 offsets) as b200_reads_t wants names and comments."""
    bs = [x if isinstance(x, bytes) else x.encode() for x in strs]
    off = np.zeros(len(bs) + 1, dtype=np.int64)
    np.cumsum([len(b) for b in bs], out=off[1:])
    return np.frombuffer(b"".join(bs) + b"\0", dtype=np.uint8), off


def numbered_names(prefix: bytes, first: int, n: int):
    """Names prefix + decimal(first + i), i < n, as (flat uint8 array, int64 offsets), vectorised (10^8 names in seconds)."""
    idx = np.arange(first, first + n, dtype=np.int64)
    nd = np.ones(n, dtype=np.int64)
    for k in range(1, 19):
        nd += idx >= 10 ** k
    ln = nd + len(prefix)
    off = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(ln, out=off[1:])
    flat = np.empty(int(off[n]) + 1, dtype=np.uint8)
    for j, ch in enumerate(prefix):
        flat[off[:-1] + j] = ch
    for d in range(1, int(nd.max()) + 1 if n else 1):          # digit d counted from the right
        m = nd >= d
        flat[off[1:][m] - d] = 48 + (idx[m] // 10 ** (d - 1)) % 10
    return flat, off


def reads_to_sam(ctxs, opt, ann_names, seqs, off, names_flat, name_off, quals=None, comments=None, n_processed=0, rg=None,
                 sub_reads=0, sam=None, want_off=True, pes0=None):
    """b200_reads_to_sam over the contexts `ctxs` (one per device).  Returns (SAM bytes as a numpy view, offsets, stats)."""
    lib = ctxs[0].lib
    n = len(off) - 1
    R = ReadsT()
    R.n_reads = n
    R.seq = seqs.ctypes.data; R.seq_off = off.ctypes.data
    R.qual = quals.ctypes.data if quals is not None else None
    R.name = names_flat.ctypes.data; R.name_off = name_off.ctypes.data
    if comments is not None:
        cf, co = comments
        R.comment = cf.ctypes.data; R.comment_off = co.ctypes.data
    a_names = (C.c_char_p * max(1, len(ann_names)))(*[s.encode() for s in ann_names])
    hs = (C.c_void_p * len(ctxs))(*[c.h for c in ctxs])
    soff = np.zeros(n + 1, dtype=np.int64) if want_off else None
    stats = (C.c_double * 8)()
    if sam is None:
        sam = np.zeros(max(1 << 16, int(off[-1] - off[0]) * 2 + 300 * n), dtype=np.uint8)
    while True:
        rc = lib.b200_reads_to_sam(hs, len(ctxs), C.byref(opt), a_names, None, C.byref(R), C.c_int64(n_processed),
                                   C.c_void_p(pes0.ctypes.data) if pes0 is not None else None,
                                   rg.encode() if rg else None, C.c_int(sub_reads), C.c_int64(len(sam)), C.c_void_p(sam.ctypes.data),
                                   C.c_void_p(soff.ctypes.data) if want_off else None, stats)
        if E_CAPACITY < rc < 0:
            sam = np.zeros(int(-rc) + 64, dtype=np.uint8)
            continue
        ctxs[0]._chk(rc, "reads_to_sam")
        return sam[:rc], soff, list(stats)


def pack_pac(contigs):
    """2-bit forward reference in the layout of the reference's .pac (bntseq.c:230: base i in byte i>>2, bits (~i&3)*2)."""
    l_pac = int(sum(len(c) for c in contigs))
    pac = np.zeros(l_pac // 4 + 1, dtype=np.uint8)
    pos = 0
    for c in contigs:                                    # contig by contig: no second copy of a 3.1 Gbp genome
        c = np.asarray(c, dtype=np.uint8)
        if c.size and int(c.max()) > 3:
            raise B200Error("bases must be nt4 codes 0..3")
        head = (-pos) % 4                                # bases completing the current byte
        for k in range(min(head, len(c))):
            i = pos + k
            pac[i >> 2] |= np.uint8(int(c[k]) << ((~i & 3) << 1))
        body = c[head:]
        nb = len(body) // 4
        if nb:
            q = body[: nb * 4].reshape(nb, 4)
            b0 = (pos + head) >> 2
            pac[b0: b0 + nb] = (q[:, 0] << 6) | (q[:, 1] << 4) | (q[:, 2] << 2) | q[:, 3]
        for k in range(nb * 4, len(body)):
            i = pos + head + k
            pac[i >> 2] |= np.uint8(int(body[k]) << ((~i & 3) << 1))
        pos += len(c)
    return pac, l_pac


class DeviceBuiltIndex:
    """Host-side view of an index built on the device by Context.build_index_device (same attributes as HostIndex)."""

    def __init__(self, ctx: "Context", contigs, names=None, pac=None, contig_lens=None):
        self.ctx = ctx
        if pac is not None:                                   # already packed (bntseq.c:230 layout), contig lengths given
            self._pac, self.l_pac = np.ascontiguousarray(pac, dtype=np.uint8), int(sum(contig_lens))
            lens = np.array(contig_lens, dtype=np.int64)
            contigs = contig_lens
        else:
            self._pac, self.l_pac = pack_pac(contigs)
            lens = np.array([len(c) for c in contigs], dtype=np.int64)
        self.ann_off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
        self.ann_len = lens.astype(np.int32)
        self.ann_alt = np.zeros(len(contigs), dtype=np.int32)
        self.names = names or [f"chr{i + 1}" for i in range(len(contigs))]
        self._bwt = self._sa = None

    def _fill_info(self):
        info = self.ctx.index_info()
        self.primary = int(info[0])
        self.L2 = [int(x) for x in info[1:6]]
        self.seq_len, self.bwt_size, self.n_sa, self.sa_intv = int(info[6]), int(info[7]), int(info[8]), int(info[9])
        self.sort_groups, self.host_ordered = int(info[11]), int(info[12])

    def _download(self):
        if self._bwt is None:
            self._bwt = np.zeros(self.bwt_size, dtype=np.uint32)
            self._sa = np.zeros(self.n_sa, dtype=np.uint64)
            self.ctx._chk(self.ctx.lib.b200_index_download(self.ctx.h, _p(self._bwt, C.c_uint32), _p(self._sa, C.c_uint64)),
                          "index_download")

    def bwt_words(self):
        self._download()
        return self._bwt

    def sa(self):
        self._download()
        return self._sa

    def pac(self):
        return self._pac

    def close(self):
        self._bwt = self._sa = None


class Context:
    """One GPU: device index + batch pipeline (mirrors process_data_t / transferIndex of the reference)."""

    def __init__(self, device: int = 0, lib: C.CDLL | None = None):
        self.lib = lib or load()
        h = self.lib.b200_ctx_create(device)
        if not h:
            raise B200Error(self.lib.b200_last_error().decode())
        self.h = C.c_void_p(h)

    def close(self):
        if self.h:
            self.lib.b200_ctx_destroy(self.h)
            self.h = None

    def trim(self):
        self._chk(self.lib.b200_ctx_trim(self.h), "trim")

    def tune(self, key: str, value: int):
        self._chk(self.lib.b200_ctx_tune(self.h, key.encode(), C.c_int64(value)), "tune")

    def microbench(self, kind: int, table_bytes: int = 0, blocks_per_sm: int = 8, iters: int = 2000, mlp: int = 1):
        """(milliseconds, units): kind 0 random 32-B sector gathers, kind 1 DPX instructions."""
        out = (C.c_double * 2)()
        self._chk(self.lib.b200_microbench(self.h, C.c_int(kind), C.c_int64(table_bytes), C.c_int(blocks_per_sm),
                                           C.c_int(iters), C.c_int(mlp), out), "microbench")
        return out[0], out[1]

    def _chk(self, rc, what):
        if rc < 0:
            raise B200Error(f"{what}: {self.lib.b200_last_error().decode()} (rc={rc})")
        return rc

    # ---------------- index
    def upload_index_structs(self, bwt_ptr, bns_ptr, pac_ptr, flags=IDX_FULL_SA):
        self._chk(self.lib.b200_index_upload(self.h, C.c_void_p(bwt_ptr), C.c_void_p(bns_ptr), C.c_void_p(pac_ptr),
                                             C.c_int(flags)), "index_upload")

    def upload_index(self, bwt_words, primary, L2, seq_len, sa, sa_intv, pac, l_pac, ann_off, ann_len, ann_alt,
                     flags=IDX_FULL_SA):
        bwt_words = np.ascontiguousarray(bwt_words, dtype=np.uint32)
        sa = np.ascontiguousarray(sa, dtype=np.uint64)
        pac = np.ascontiguousarray(pac, dtype=np.uint8)
        assert len(pac) >= l_pac // 4 + 1
        ann_off = np.ascontiguousarray(ann_off, dtype=np.int64)
        ann_len = np.ascontiguousarray(ann_len, dtype=np.int32)
        ann_alt = np.ascontiguousarray(ann_alt, dtype=np.int32)
        l2 = (C.c_uint64 * 5)(*[int(x) for x in L2])
        self._chk(self.lib.b200_index_upload_raw(
            self.h, _p(bwt_words, C.c_uint32), C.c_uint64(len(bwt_words)), C.c_uint64(primary), l2,
            C.c_uint64(seq_len), _p(sa, C.c_uint64), C.c_uint64(len(sa)), C.c_int(sa_intv), _p(pac, C.c_uint8),
            C.c_int64(l_pac), C.c_int(len(ann_off)), _p(ann_off, C.c_int64), _p(ann_len, C.c_int32),
            _p(ann_alt, C.c_int32), C.c_int(flags)), "index_upload_raw")

    def upload_host_index(self, hidx: "HostIndex", flags=IDX_FULL_SA):
        self._chk(self.lib.b200_index_upload_host(self.h, hidx.h, C.c_int(len(hidx.ann_off)), _p(hidx.ann_off, C.c_int64),
                                                  _p(hidx.ann_len, C.c_int32), _p(hidx.ann_alt, C.c_int32),
                                                  C.c_int(flags)), "index_upload_host")

    def build_index_device(self, contigs, flags=IDX_FULL_SA, names=None):
        """Index the genome ON the device (b200_index_build_device) and leave it resident.  Returns a DeviceBuiltIndex
        with the attributes of HostIndex (arrays are downloaded on demand)."""
        d = DeviceBuiltIndex(self, contigs, names)
        self._chk(self.lib.b200_index_build_device(
            self.h, _p(d._pac, C.c_uint8), C.c_int64(d.l_pac), C.c_int(len(d.ann_off)), _p(d.ann_off, C.c_int64),
            _p(d.ann_len, C.c_int32), _p(d.ann_alt, C.c_int32), C.c_int(flags)), "index_build_device")
        d._fill_info()
        return d

    def build_index_device_pac(self, pac, contig_lens, flags=IDX_FULL_SA, names=None):
        """The same from the 2-bit packed forward reference (`.pac` layout) and the contig lengths."""
        d = DeviceBuiltIndex(self, None, names, pac=pac, contig_lens=contig_lens)
        self._chk(self.lib.b200_index_build_device(
            self.h, _p(d._pac, C.c_uint8), C.c_int64(d.l_pac), C.c_int(len(d.ann_off)), _p(d.ann_off, C.c_int64),
            _p(d.ann_len, C.c_int32), _p(d.ann_alt, C.c_int32), C.c_int(flags)), "index_build_device")
        d._fill_info()
        return d

    def save_index(self, path: str, with_full_sa: bool = False):
        self._chk(self.lib.b200_index_save(self.h, path.encode(), C.c_int(1 if with_full_sa else 0)), "index_save")

    def load_index(self, path: str, flags=IDX_FULL_SA):
        self._chk(self.lib.b200_index_load(self.h, path.encode(), C.c_int(flags)), "index_load")

    def replicate_index_from(self, src: "Context"):
        """b200_index_replicate: borrow (same device) or peer-copy (other device) src's index."""
        self._chk(self.lib.b200_index_replicate(self.h, src.h), "index_replicate")
        self._index_src = src                                  # keep the owner alive

    def index_info(self):
        info = np.zeros(13, dtype=np.uint64)
        self._chk(self.lib.b200_index_info(self.h, C.c_void_p(info.ctypes.data)), "index_info")
        return info

    # ---------------- primitives
    def occ4(self, k):
        k = np.ascontiguousarray(k, dtype=np.uint64)
        out = np.zeros((len(k), 4), dtype=np.uint64)
        self._chk(self.lib.b200_occ4_batch(self.h, C.c_int64(len(k)), _p(k, C.c_uint64), _p(out, C.c_uint64)), "occ4")
        return out

    def occ2x4(self, k, l):
        k = np.ascontiguousarray(k, dtype=np.uint64)
        l = np.ascontiguousarray(l, dtype=np.uint64)
        ok = np.zeros((len(k), 4), dtype=np.uint64)
        ol = np.zeros((len(k), 4), dtype=np.uint64)
        self._chk(self.lib.b200_2occ4_batch(self.h, C.c_int64(len(k)), _p(k, C.c_uint64), _p(l, C.c_uint64),
                                            _p(ok, C.c_uint64), _p(ol, C.c_uint64)), "2occ4")
        return ok, ol

    def extend(self, ik, is_back):
        ik = np.ascontiguousarray(ik, dtype=np.uint64)
        out = np.zeros((len(ik), 4, 3), dtype=np.uint64)
        self._chk(self.lib.b200_extend_batch(self.h, C.c_int64(len(ik)), _p(ik, C.c_uint64), C.c_int(is_back),
                                             _p(out, C.c_uint64)), "extend")
        return out

    def sa_lookup(self, k, use_full=True):
        k = np.ascontiguousarray(k, dtype=np.uint64)
        out = np.zeros(len(k), dtype=np.uint64)
        self._chk(self.lib.b200_sa_batch(self.h, C.c_int64(len(k)), _p(k, C.c_uint64), _p(out, C.c_uint64),
                                         C.c_int(1 if use_full else 0)), "sa")
        return out

    def smem1a(self, seqs, off, x, min_intv, max_intv=0, cap=512):
        n = len(off) - 1
        x = np.ascontiguousarray(x, dtype=np.int32)
        min_intv = np.ascontiguousarray(min_intv, dtype=np.int32)
        out = np.zeros((n, cap), dtype=INTV_DTYPE)
        cnt = np.zeros(n, dtype=np.int32)
        ret = np.zeros(n, dtype=np.int32)
        self._chk(self.lib.b200_smem1a_batch(self.h, C.c_int64(n), _p(seqs, C.c_uint8), _p(off, C.c_int64),
                                             _p(x, C.c_int32), _p(min_intv, C.c_int32), C.c_uint64(max_intv),
                                             C.c_int(cap), C.c_void_p(out.ctypes.data), _p(cnt, C.c_int32),
                                             _p(ret, C.c_int32)), "smem1a")
        return out, cnt, ret

    def ksw_extend2(self, qlen, qoff, q, tlen, toff, t, opt: MemOpt, w, end_bonus, zdrop, h0):
        n = len(qlen)
        out = np.zeros((n, 6), dtype=np.int32)
        self._chk(self.lib.b200_ksw_extend2_batch(
            self.h, C.c_int64(n), _p(qlen, C.c_int32), _p(qoff, C.c_int64), _p(q, C.c_uint8), _p(tlen, C.c_int32),
            _p(toff, C.c_int64), _p(t, C.c_uint8), opt.mat, opt.o_del, opt.e_del, opt.o_ins, opt.e_ins,
            _p(w, C.c_int32), C.c_int(end_bonus), C.c_int(zdrop), _p(h0, C.c_int32), _p(out, C.c_int32)), "ksw_extend2")
        return out

    def ksw_align2(self, qlen, qoff, q, tlen, toff, t, opt: MemOpt, xtra):
        n = len(qlen)
        out = np.zeros((n, 7), dtype=np.int32)
        xtra = np.ascontiguousarray(xtra, dtype=np.int32)
        self._chk(self.lib.b200_ksw_align2_batch(
            self.h, C.c_int64(n), _p(qlen, C.c_int32), _p(qoff, C.c_int64), _p(q, C.c_uint8), _p(tlen, C.c_int32),
            _p(toff, C.c_int64), _p(t, C.c_uint8), opt.mat, opt.o_del, opt.e_del, opt.o_ins, opt.e_ins,
            _p(xtra, C.c_int32), _p(out, C.c_int32)), "ksw_align2")
        return out

    # ---------------- stage seams
    def collect_intv(self, opt: MemOpt, seqs, off, cap_per_read=256):
        n = len(off) - 1
        cap = max(1, n * cap_per_read)
        out = np.zeros(cap, dtype=INTV_DTYPE)
        cnt = np.zeros(max(n, 1), dtype=np.int32)
        tot = self._chk(self.lib.b200_collect_intv_batch(self.h, C.byref(opt), C.c_int(n), _p(seqs, C.c_uint8),
                                                         _p(off, C.c_int64), C.c_int64(cap),
                                                         C.c_void_p(out.ctypes.data), _p(cnt, C.c_int32)), "collect_intv")
        return out[:tot].copy(), cnt[:n]

    def chains(self, opt: MemOpt, seqs, off, stage=1, cap_chn_per_read=64, cap_seed_per_read=1024):
        n = len(off) - 1
        cc, cs = max(1, n * cap_chn_per_read), max(1, n * cap_seed_per_read)
        chn = np.zeros((cc, 8), dtype=np.int64)
        sd = np.zeros((cs, 4), dtype=np.int64)
        nch = np.zeros(max(n, 1), dtype=np.int32)
        tot = np.zeros(2, dtype=np.int64)
        self._chk(self.lib.b200_chain_batch(self.h, C.byref(opt), C.c_int(stage), C.c_int(n), _p(seqs, C.c_uint8),
                                            _p(off, C.c_int64), C.c_int64(cc), _p(chn, C.c_int64), _p(nch, C.c_int32),
                                            C.c_int64(cs), _p(sd, C.c_int64), _p(tot, C.c_int64)), "chains")
        return chn[:tot[0]].copy(), nch[:n], sd[:tot[1]].copy()

    # ---------------- the hot path: worker1 over a batch
    def align(self, opt: MemOpt, seqs, off, cap_per_read=8):
        """mem_align1_core for every read; host buffers in/out (copies included)."""
        n = len(off) - 1
        cap = max(64, n * cap_per_read)
        cnt = np.zeros(max(n, 1), dtype=np.int32)
        while True:
            out = np.zeros(cap, dtype=ALNREG_DTYPE)
            tot = self.lib.b200_align_batch(self.h, C.byref(opt), C.c_int(n), _p(seqs, C.c_uint8), _p(off, C.c_int64),
                                            C.c_int64(cap), C.c_void_p(out.ctypes.data), _p(cnt, C.c_int32))
            if tot >= 0:
                return out[:tot], cnt[:n]
            if tot > E_CAPACITY and -tot > cap:
                cap = int(-tot)
                continue
            self._chk(tot, "align")

    def reg2aln(self, opt: MemOpt, seqs, off, regs, cnt):
        """mem_reg2aln for every region: (b200_aln_t records, CIGAR words, MD bytes)."""
        n = len(off) - 1
        regs = np.ascontiguousarray(regs); cnt = np.ascontiguousarray(cnt, dtype=np.int32)
        aln = np.zeros(max(1, len(regs)), dtype=ALN_DTYPE)
        tot = (C.c_int64 * 2)()
        cap_c, cap_m = 8 * len(regs) + int(off[-1] - off[0]) // 8 + 64, 24 * len(regs) + int(off[-1] - off[0]) // 4 + 64
        while True:
            cig = np.zeros(cap_c, dtype=np.uint32); md = np.zeros(cap_m, dtype=np.uint8)
            rc = self.lib.b200_reg2aln_batch(self.h, C.byref(opt), C.c_int(n), _p(seqs, C.c_uint8), _p(off, C.c_int64),
                                             C.c_void_p(regs.ctypes.data), _p(cnt, C.c_int32), C.c_void_p(aln.ctypes.data),
                                             C.c_int64(cap_c), _p(cig, C.c_uint32), C.c_int64(cap_m), C.c_void_p(md.ctypes.data), tot)
            if rc == -2:
                cap_c, cap_m = int(tot[0]) + 64, int(tot[1]) + 64
                continue
            self._chk(rc, "reg2aln")
            return aln[:rc], cig[: int(tot[0])], md[: int(tot[1])]

    def reg2sam_se(self, opt: MemOpt, seqs, off, regs, cnt, n_processed=0):
        """worker2's single-end half before the SAM text (b200_reg2sam_se_batch): primary marking, record selection, MAPQ,
        XA choice and mem_reg2aln -> (regions as updated, b200_sel_t, b200_aln_t, CIGAR words, MD bytes)."""
        from .abi import SEL_DTYPE
        n = len(off) - 1
        regs0 = np.array(regs, dtype=ALNREG_DTYPE, copy=True); cnt = np.ascontiguousarray(cnt, dtype=np.int32)
        regs = regs0.copy()
        aln = np.zeros(max(1, len(regs)), dtype=ALN_DTYPE)
        sel = np.zeros(max(1, len(regs)), dtype=SEL_DTYPE)
        tot = (C.c_int64 * 2)()
        cap_c, cap_m = 8 * len(regs) + int(off[-1] - off[0]) // 8 + 64, 24 * len(regs) + int(off[-1] - off[0]) // 4 + 64
        while True:
            cig = np.zeros(cap_c, dtype=np.uint32); md = np.zeros(cap_m, dtype=np.uint8)
            regs[...] = regs0
            rc = self.lib.b200_reg2sam_se_batch(self.h, C.byref(opt), C.c_int(n), C.c_int64(n_processed), _p(seqs, C.c_uint8),
                                                _p(off, C.c_int64), C.c_void_p(regs.ctypes.data), _p(cnt, C.c_int32),
                                                C.c_void_p(sel.ctypes.data), C.c_void_p(aln.ctypes.data), C.c_int64(cap_c),
                                                _p(cig, C.c_uint32), C.c_int64(cap_m), C.c_void_p(md.ctypes.data), tot)
            if rc == -2:
                cap_c, cap_m = int(tot[0]) + 64, int(tot[1]) + 64
                continue
            self._chk(rc, "reg2sam_se")
            return regs[:rc], sel[:rc], aln[:rc], cig[: int(tot[0])], md[: int(tot[1])]

    def align_sam_se(self, opt: MemOpt, ann_names, names, seqs, off, n_processed=0, quals=None, n_threads=8, sub_reads=0):
        """b200_align_sam_se_batch: reads -> SAM text (bytes) and the n + 1 offsets of the reads' records, pipelined."""
        n = len(off) - 1
        a_names = (C.c_char_p * max(1, len(ann_names)))(*[s.encode() for s in ann_names])
        r_names = (C.c_char_p * max(1, n))(*[s.encode() for s in names])
        soff = np.zeros(n + 1, dtype=np.int64)
        cap = 1 << 16
        while True:
            sam = np.zeros(cap, dtype=np.uint8)
            rc = self.lib.b200_align_sam_se_batch(self.h, C.byref(opt), a_names, None, C.c_int(n), C.c_int64(n_processed), r_names, None,
                                                  _p(seqs, C.c_uint8), None if quals is None else _p(quals, C.c_uint8), _p(off, C.c_int64),
                                                  None, C.c_int(n_threads), C.c_int(sub_reads), C.c_int64(cap), C.c_void_p(sam.ctypes.data),
                                                  _p(soff, C.c_int64))
            if E_CAPACITY < rc < 0:
                cap = int(-rc) + 64
                continue
            self._chk(rc, "align_sam_se")
            return sam[:rc].tobytes(), soff

    def batch_reg2sam_se(self, opt: MemOpt, cnt, n_processed=0):
        """b200_batch_reg2sam_se: worker2 on the batch this context has just aligned (nothing uploaded)."""
        from .abi import SEL_DTYPE
        cnt = np.ascontiguousarray(cnt, dtype=np.int32)
        nr = int(cnt.astype(np.int64).sum())
        regs = np.zeros(max(1, nr), dtype=ALNREG_DTYPE)
        aln = np.zeros(max(1, nr), dtype=ALN_DTYPE); sel = np.zeros(max(1, nr), dtype=SEL_DTYPE)
        tot = (C.c_int64 * 2)()
        cap_c, cap_m = 8 * nr + 64, 24 * nr + 64
        while True:
            cig = np.zeros(cap_c, dtype=np.uint32); md = np.zeros(cap_m, dtype=np.uint8)
            rc = self.lib.b200_batch_reg2sam_se(self.h, C.byref(opt), C.c_int64(n_processed), _p(cnt, C.c_int32),
                                                C.c_void_p(regs.ctypes.data), C.c_void_p(sel.ctypes.data), C.c_void_p(aln.ctypes.data),
                                                C.c_int64(cap_c), _p(cig, C.c_uint32), C.c_int64(cap_m), C.c_void_p(md.ctypes.data), tot)
            if rc == -2:
                cap_c, cap_m = int(tot[0]) + 64, int(tot[1]) + 64
                continue
            self._chk(rc, "batch_reg2sam_se")
            return regs[:rc], sel[:rc], aln[:rc], cig[: int(tot[0])], md[: int(tot[1])]

    def sam_se(self, opt: MemOpt, ann_names, names, seqs, off, cnt, sel, aln, cig, md, quals=None, comments=None, rg_id=None,
               n_threads=8):
        """b200_sam_se_format: SAM text (bytes) and the n + 1 offsets of the reads' records."""
        n = len(off) - 1
        cnt = np.ascontiguousarray(cnt, dtype=np.int32)
        sel = np.ascontiguousarray(sel); aln = np.ascontiguousarray(aln)
        cig = np.ascontiguousarray(cig, dtype=np.uint32); md = np.ascontiguousarray(md, dtype=np.uint8)
        a_names = (C.c_char_p * max(1, len(ann_names)))(*[s.encode() for s in ann_names])
        r_names = (C.c_char_p * max(1, n))(*[s.encode() for s in names])
        r_comm = None if comments is None else (C.c_char_p * max(1, n))(*[None if s is None else s.encode() for s in comments])
        soff = np.zeros(n + 1, dtype=np.int64)
        cap = 1 << 16
        while True:
            sam = np.zeros(cap, dtype=np.uint8)
            rc = self.lib.b200_sam_se_format(C.byref(opt), a_names, None, C.c_int(n), r_names, r_comm, _p(seqs, C.c_uint8),
                                             None if quals is None else _p(quals, C.c_uint8), _p(off, C.c_int64), _p(cnt, C.c_int32),
                                             C.c_void_p(sel.ctypes.data), C.c_void_p(aln.ctypes.data), _p(cig, C.c_uint32),
                                             C.c_void_p(md.ctypes.data), None if rg_id is None else rg_id.encode(), C.c_int(n_threads),
                                             C.c_int64(cap), C.c_void_p(sam.ctypes.data), _p(soff, C.c_int64))
            if rc < 0:
                cap = int(-rc) + 64
                continue
            return sam[:rc].tobytes(), soff

    def batch_upload(self, seqs, off):
        n = len(off) - 1
        self._chk(self.lib.b200_batch_upload(self.h, C.c_int(n), _p(seqs, C.c_uint8), _p(off, C.c_int64)), "batch_upload")

    def batch_run(self, opt: MemOpt, want_stage_ms=False):
        ms = (C.c_float * 6)()
        self._chk(self.lib.b200_batch_run(self.h, C.byref(opt), ms if want_stage_ms else None), "batch_run")
        return list(ms) if want_stage_ms else None

    def batch_download(self, n_reads, cap):
        out = np.zeros(cap, dtype=ALNREG_DTYPE)
        cnt = np.zeros(max(n_reads, 1), dtype=np.int32)
        tot = self.lib.b200_batch_download(self.h, C.c_int64(cap), C.c_void_p(out.ctypes.data), _p(cnt, C.c_int32))
        if tot <= E_CAPACITY:
            self._chk(tot, "batch_download")
        if tot < 0:
            return None, int(-tot)
        return out[:tot], cnt[:n_reads]

    def batch_set_regs(self, regs, cnt):
        regs = np.ascontiguousarray(regs); cnt = np.ascontiguousarray(cnt, dtype=np.int32)
        self._chk(self.lib.b200_batch_set_regs(self.h, C.c_void_p(regs.ctypes.data), _p(cnt, C.c_int32)), "batch_set_regs")

    def batch_rescue(self, opt: MemOpt, pes):
        """pes: 4 x mem_pestat_t as a numpy array of PESTAT_DTYPE.  Returns (ksw_align2 calls, DP cells, device ms)."""
        pes = np.ascontiguousarray(pes)
        st = (C.c_int64 * 2)(); ms = C.c_float(0)
        self._chk(self.lib.b200_batch_rescue(self.h, C.byref(opt), C.c_void_p(pes.ctypes.data), st, C.byref(ms)), "batch_rescue")
        return int(st[0]), int(st[1]), float(ms.value)

    def counters(self):
        out = np.zeros(16, dtype=np.uint64)
        self.lib.b200_batch_counters(self.h, _p(out, C.c_uint64))
        names = ["extend", "two_block", "sa", "lf_steps", "ksw", "cells", "seeds", "chains", "regs", "launches",
                 "seed_sw", "seed_sw_cells", "ext_bin0", "ext_bin1", "ext_bin2", "dp8_jobs"]
        return {k: int(out[i]) for i, k in enumerate(names)}
