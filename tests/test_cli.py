"""The reference's own `bwa` program linked against libbwamem_b200.so (scripts/build_bwa.sh: main.c, fastmap.c with the
three-line seam at fastmap.c:35,309,311, bwa.c, bntseq.c, bwtindex.c ... compiled from /root/reference where they lie):
`bwa mem` with real option parsing (fastmap.c:77-175, -x presets :240-268, update_a :43-57), FASTQ input through bseq_read
and batches of -K bases, against `bwa_cpu`, the same program with the reference's own CPU mem_process_seqs.  The binaries
are built by __graft_entry__.build() where the reference is present and travel to the GPU box."""
import os
import subprocess

import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import synth

BWA = os.path.join(cm.ROOT, "cli", "_build", "bwa")
BWA_CPU = os.path.join(cm.ROOT, "cli", "_build", "bwa_cpu")
have_cli = os.path.exists(BWA) and os.path.exists(BWA_CPU)
needs_cli = pytest.mark.skipif(not have_cli, reason="cli/_build not built (scripts/build_bwa.sh needs the reference sources)")


def _body(sam: bytes) -> bytes:
    return b"\n".join(l for l in sam.split(b"\n") if not l.startswith(b"@PG"))


def _run(exe, args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([exe] + args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=e, timeout=900)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    return r.stdout


@pytest.fixture(scope="module")
def workdir(tmp_path_factory):
    d = tmp_path_factory.mktemp("cli")
    contigs = synth.make_genome([400_000, 250_000, 90_000], seed=77, repeat_frac=0.02)
    fa = str(d / "ref.fa")
    synth.write_fasta(fa, contigs)
    _run(BWA_CPU, ["index", fa])
    n = 6000
    reads, _ = synth.simulate_reads(contigs, n, seed=5, indel_rate=0.002, sub_rate=0.015)
    seqs = np.ascontiguousarray(reads.reshape(-1)); off = np.arange(n + 1, dtype=np.int64) * 150
    rng = np.random.default_rng(1)
    quals = rng.integers(35, 74, size=len(seqs), dtype=np.uint8)
    synth.write_fastq(str(d / "se.fq"), seqs, off, [f"r{i}" for i in range(n)], quals, [f"BC:Z:x{i % 7}" for i in range(n)])
    pairs, _ = synth.simulate_pairs(contigs, 3000, seed=6, damaged_frac=0.1, damage="periodic")
    p = np.ascontiguousarray(pairs.reshape(-1)); poff = np.arange(6001, dtype=np.int64) * 150
    synth.write_fastq(str(d / "pe_1.fq"), np.ascontiguousarray(pairs[0::2].reshape(-1)), np.arange(3001, dtype=np.int64) * 150, [f"p{i}" for i in range(3000)])
    synth.write_fastq(str(d / "pe_2.fq"), np.ascontiguousarray(pairs[1::2].reshape(-1)), np.arange(3001, dtype=np.int64) * 150, [f"p{i}" for i in range(3000)])
    # interleaved file with a few unpaired reads in between (smart pairing, -p)
    names = [f"p{i // 2}" for i in range(6000)]
    names[100] = "solo_a"; names[101] = "solo_b"; names[2000] = "solo_c"; names[2001] = "solo_d"
    synth.write_fastq(str(d / "inter.fq"), p, poff, names)
    return d, fa


@needs_cli
def test_cli_cpu_program_matches_library_call(workdir):
    """Sanity of the recipe (no GPU): `bwa_cpu mem` = the reference's mem_process_seqs on the same batch."""
    d, fa = workdir
    out = _run(BWA_CPU, ["mem", "-t", "4", "-K", "100000000", fa, str(d / "se.fq")])
    body = [l for l in out.split(b"\n") if l and not l.startswith(b"@")]
    assert len(body) >= 6000 and body[0].startswith(b"r0\t")
    assert out.startswith(b"@SQ\tSN:chr1\tLN:400000")


@needs_cli
@pytest.mark.gpu
@pytest.mark.parametrize("args", [
    ["-t", "4", "-K", "100000000"],
    ["-t", "4", "-K", "300000", "-C", "-R", "@RG\\tID:grp\\tSM:s1"],              # several batches, comments copied, read group
    ["-t", "2", "-K", "100000000", "-A", "2", "-Y", "-M", "-a"],                    # update_a scaling, soft clips, -M, all hits
    ["-t", "4", "-K", "100000000", "-x", "intractg", "-T", "40", "-5"],
])
def test_cli_single_end(workdir, args):
    d, fa = workdir
    exp = _run(BWA_CPU, ["mem"] + args + [fa, str(d / "se.fq")])
    got = _run(BWA, ["mem"] + args + [fa, str(d / "se.fq")])
    assert _body(got) == _body(exp)


@needs_cli
@pytest.mark.gpu
@pytest.mark.parametrize("args,files", [
    (["-t", "4", "-K", "100000000"], ["pe_1.fq", "pe_2.fq"]),
    (["-t", "4", "-K", "400000", "-5", "-Y"], ["pe_1.fq", "pe_2.fq"]),               # several mem_pestat batches; Hi-C style flags
    (["-t", "4", "-K", "100000000", "-p"], ["inter.fq"]),                           # smart pairing with unpaired reads in the file
    (["-t", "4", "-K", "100000000", "-I", "500,50,800,200"], ["pe_1.fq", "pe_2.fq"]),   # insert-size statistics given (pes0)
])
def test_cli_paired_end(workdir, args, files):
    d, fa = workdir
    f = [str(d / x) for x in files]
    exp = _run(BWA_CPU, ["mem"] + args + [fa] + f)
    got = _run(BWA, ["mem"] + args + [fa] + f)
    assert _body(got) == _body(exp)


@needs_cli
@pytest.mark.gpu
def test_cli_two_devices(workdir):
    """B200_DEVICES=2: every batch cut over two contexts (two GPUs when the box has them, else the call fails over to one)."""
    d, fa = workdir
    args = ["mem", "-t", "4", "-K", "100000000", fa, str(d / "pe_1.fq"), str(d / "pe_2.fq")]
    exp = _run(BWA_CPU, args)
    import torch
    nd = 2 if torch.cuda.device_count() >= 2 else 1
    got = _run(BWA, args, env={"B200_DEVICES": str(nd)})
    assert _body(got) == _body(exp)
