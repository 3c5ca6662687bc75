"""b200_reads_to_sam (csrc/reads2sam.cu + csrc/samdev.cu): reads in host memory -> SAM text in host memory with worker1,
worker2 AND the SAM text on the device, over one or several contexts, against the reference's mem_process_seqs
(bwamem.c:1215) byte for byte: qualities, comments, RG, ragged reads, several sub-batch sizes, several contexts."""
import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import abi, lib as L, synth

BACKENDS = [pytest.param("emu", id="emu"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]
needs_ref = pytest.mark.skipif(not cm.have_ref(), reason="oracle/_ref not built")


def test_numbered_names():
    for first, n in ((0, 1), (0, 1234), (95, 20), (999_990, 25)):
        flat, off = L.numbered_names(b"r", first, n)
        exp = [b"r%d" % (first + i) for i in range(n)]
        assert [bytes(flat[off[i]:off[i + 1]]) for i in range(n)] == exp


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname", ["small", "dup"])
def test_reads_to_sam_matches_reference(backend, gname):
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    opt = abi.default_opt()
    opt.n_threads = 4
    n = 300 if backend == "emu" else 30000
    seqs, off = cm.reads_for(gname, n, seed=171, indel=0.004, sub=0.015)
    rs, ro = cm.ragged_reads(gname)                          # empty, short, all-N, chimeric, long reads in the middle of the batch
    seqs = np.concatenate([seqs[: off[n // 2]], rs, seqs[off[n // 2]:]])
    lens = np.concatenate([np.diff(off[: n // 2 + 1]), np.diff(ro), np.diff(off[n // 2:])])
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    n = len(off) - 1
    rng = np.random.default_rng(3)
    quals = rng.integers(35, 74, size=len(seqs), dtype=np.uint8)
    names = [f"read_{i * 7919 % 1000003}" for i in range(n)]
    comments = [f"BC:Z:{i % 97}" if i % 3 else None for i in range(n)]
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, names=names, quals=quals, comments=comments, n_processed=4242)
    nf, no = L.flat_strings(names)
    cf = L.flat_strings([c or "" for c in comments])
    for sub in (0, 128 if backend == "emu" else 7000):
        sam, soff, stats = L.reads_to_sam([ctx], opt, idx.anns()[3], seqs, off, nf, no, quals=quals, comments=cf, n_processed=4242,
                                          sub_reads=sub)
        got = sam.tobytes()
        if got != exp:
            for x, y in zip(exp.split(b"\n"), got.split(b"\n")):
                assert x == y
        assert got == exp
        assert soff[0] == 0 and soff[-1] == len(got) and got[soff[n // 2]:].startswith(names[n // 2].encode() + b"\t")
    # ASCII input, a read group, too small a buffer (-(needed), then the same text), and two contexts sharing the sub-batches
    ctx2 = L.Context(0, ctx.lib)
    ctx2.replicate_index_from(ctx)                              # same device: the index is borrowed, not copied
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, names=names, n_processed=0, rg="grp1")
    small = np.zeros(1000, dtype=np.uint8)
    sam, soff, stats = L.reads_to_sam([ctx, ctx2], opt, idx.anns()[3], synth.to_ascii(seqs), off, nf, no, rg="grp1",
                                      sub_reads=64 if backend == "emu" else 5000, sam=small)
    assert sam.tobytes() == exp and stats[4] >= 3
    ctx2.close()


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_reads_to_sam_nothing_aligned(backend):
    """Reads that hit nothing: every one gets the unmapped record (bwamem.c:1043-1047), still through the device formatter."""
    idx, ctx = cm.ref_index("tiny"), cm.context(backend, "tiny")
    opt = abi.default_opt()
    rng = np.random.default_rng(11)
    seqs = rng.integers(0, 4, size=40 * 60, dtype=np.uint8)
    off = np.arange(41, dtype=np.int64) * 60
    seqs[::7] = 4
    names = [f"u{i}" for i in range(40)]
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, names=names)
    nf, no = L.flat_strings(names)
    sam, _, _ = L.reads_to_sam([ctx], opt, idx.anns()[3], seqs, off, nf, no)
    assert sam.tobytes() == exp


def _pairs(gname, npairs, seed, damage, damaged_frac=0.1):
    reads, _ = synth.simulate_pairs(cm.genome(gname), npairs, seed=seed, damaged_frac=damaged_frac, damage=damage)
    seqs = np.ascontiguousarray(reads.reshape(-1))
    off = np.arange(0, 2 * npairs + 1, dtype=np.int64) * reads.shape[1]
    names = [f"p{i // 2}" for i in range(2 * npairs)]
    return seqs, off, names


def _cmp_sam(got, exp):
    if got != exp:
        for x, y in zip(exp.split(b"\n"), got.split(b"\n")):
            assert x == y
    assert got == exp


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("gname,damage", [("small", "window"), ("small", "periodic"), ("dup", "periodic"), ("rep", "window")])
def test_pairs_to_sam_matches_reference(backend, gname, damage):
    """Interleaved pairs through b200_reads_to_sam with MEM_F_PE: worker1, the batch's mem_pestat (host reduction over the
    device-made per-pair summaries), mate rescue, mem_pair / MAPQ / flags (finalize_pe.cuh), mem_reg2aln and the SAM text with
    mate fields on the device -- against the reference's mem_process_seqs (bwamem.c:1215; mem_sam_pe bwamem_pair.c:276)."""
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    opt = abi.default_opt()
    opt.flag |= 0x2
    opt.n_threads = 4
    npairs = 250 if backend == "emu" else 10000
    seqs, off, names = _pairs(gname, npairs, 181, damage)
    quals = np.random.default_rng(9).integers(35, 74, size=len(seqs), dtype=np.uint8)
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, names=names, quals=quals, n_processed=20)
    nf, no = L.flat_strings(names)
    for sub in (0, 100 if backend == "emu" else 3000):
        sam, soff, _ = L.reads_to_sam([ctx], opt, idx.anns()[3], seqs, off, nf, no, quals=quals, n_processed=20, sub_reads=sub)
        _cmp_sam(sam.tobytes(), exp)
        assert soff[-1] == len(exp)
    assert b"\t99\t" in exp or b"\t83\t" in exp           # proper pairs exist
    # option variants: -5 -Y (Hi-C usage), -a, -P (no pairing), -S (no rescue), given insert-size statistics
    for flag in (0x800 | 0x200, 0x8, 0x4, 0x20):
        o = abi.default_opt()
        o.flag |= 0x2 | flag
        o.n_threads = 4
        exp = idx.process_seqs(o, synth.to_ascii(seqs), off, names=names, n_processed=0)
        sam, _, _ = L.reads_to_sam([ctx], o, idx.anns()[3], seqs, off, nf, no, sub_reads=sub)
        _cmp_sam(sam.tobytes(), exp)
    pes = np.zeros(4, dtype=abi.PESTAT_DTYPE)
    pes["failed"] = 1
    pes[1] = (300, 700, 0, 500.0, 50.0)
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, names=names, n_processed=0, pes0=pes)
    sam, _, _ = L.reads_to_sam([ctx], opt, idx.anns()[3], seqs, off, nf, no, sub_reads=sub, pes0=pes)
    _cmp_sam(sam.tobytes(), exp)


@needs_ref
@pytest.mark.parametrize("backend", BACKENDS)
def test_pairs_to_sam_two_contexts_and_few_pairs(backend):
    """Two contexts share the sub-batches of a paired batch (phase A regions stay where they were made); a batch with too
    few pairs for any orientation (every mem_pestat orientation fails: no rescue windows, no pairing)."""
    gname = "small"
    idx, ctx = cm.ref_index(gname), cm.context(backend, gname)
    ctx2 = L.Context(0, ctx.lib)
    ctx2.upload_host_index(ctx._host_index)
    opt = abi.default_opt()
    opt.flag |= 0x2
    opt.n_threads = 4
    npairs = 301 if backend == "emu" else 9001
    seqs, off, names = _pairs(gname, npairs, 183, "periodic", 0.2)
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, names=names, n_processed=64)
    nf, no = L.flat_strings(names)
    sam, _, st = L.reads_to_sam([ctx, ctx2], opt, idx.anns()[3], seqs, off, nf, no, n_processed=64, sub_reads=90 if backend == "emu" else 2500)
    _cmp_sam(sam.tobytes(), exp)
    assert st[4] >= 4
    seqs, off, names = _pairs(gname, 6, 185, "window")
    exp = idx.process_seqs(opt, synth.to_ascii(seqs), off, names=names)
    nf, no = L.flat_strings(names)
    sam, _, _ = L.reads_to_sam([ctx, ctx2], opt, idx.anns()[3], seqs, off, nf, no)
    _cmp_sam(sam.tobytes(), exp)
    ctx2.close()
