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
    ctx2.upload_host_index(ctx._host_index)
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
