"""b200_fastq_* (csrc/fastq.cpp): the input side.  Parsing and batch boundaries against a straightforward Python
restatement of kseq.h's record grammar, trim_readno (bwa.c:64-68) and bseq_read's batch rule (bwa.c:191-224): plain, gzip
and BGZF files, two mate files, multi-line FASTA, CRLF line ends, records without qualities, odd batch sizes.  The GPU test
runs FASTQ files -> SAM through the reader + b200_reads_to_sam against the reference program (tests/test_cli.py's bwa_cpu)."""
import gzip
import os
import struct
import zlib

import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import abi, lib as L, synth


def _py_records(text: bytes):
    """kseq.h grammar: (name, comment, seq, qual or None) per record."""
    lines = text.replace(b"\r\n", b"\n").split(b"\n")
    recs, i = [], 0
    while i < len(lines):
        ln = lines[i]
        if not ln or ln[:1] not in (b"@", b">"):
            i += 1
            continue
        hdr = ln[1:]
        sp = min([k for k in (hdr.find(b" "), hdr.find(b"\t")) if k >= 0], default=-1)
        name, comment = (hdr, b"") if sp < 0 else (hdr[:sp], hdr[sp + 1:])
        if len(name) > 2 and name[-2:-1] == b"/" and name[-1:].isdigit():
            name = name[:-2]
        i += 1
        seq = b""
        while i < len(lines) and lines[i][:1] not in (b"+", b">", b"@"):
            seq += lines[i]; i += 1
        qual = None
        if i < len(lines) and lines[i][:1] == b"+":
            i += 1
            qual = b""
            while i < len(lines) and len(qual) < len(seq):
                qual += lines[i]; i += 1
        recs.append((name, comment, seq, qual))
    return recs


def _py_batches(recs_per_file, chunk):
    """bseq_read: alternate the files' records, stop when size >= chunk on an even count."""
    if len(recs_per_file) == 2:
        m = min(len(recs_per_file[0]), len(recs_per_file[1]))
        order = [r for k in range(m) for r in (recs_per_file[0][k], recs_per_file[1][k])]
    else:
        order = recs_per_file[0]
    out, cur, size = [], [], 0
    for r in order:
        cur.append(r); size += len(r[2])
        if size >= chunk and len(cur) % 2 == 0:
            out.append(cur); cur, size = [], 0
    if cur:
        out.append(cur)
    return out


def _bgzf(data: bytes, block=40000) -> bytes:
    out = b""
    for b in range(0, len(data), block):
        raw = data[b:b + block]
        c = zlib.compressobj(6, zlib.DEFLATED, -15)
        d = c.compress(raw) + c.flush()
        bsize = len(d) + 25
        out += struct.pack("<4BI2BH2BHH", 31, 139, 8, 4, 0, 0, 255, 6, 66, 67, 2, bsize) + d + struct.pack("<II", zlib.crc32(raw), len(raw))
    return out + bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")     # the BGZF end-of-file block


def _check(reader, expected_batches):
    for exp in expected_batches:
        b = reader.read(CHUNK)
        assert b is not None and b["n"] == len(exp)
        nf, no = b["names"]
        for i, (name, comment, seq, qual) in enumerate(exp):
            assert bytes(nf[no[i]:no[i + 1]]) == name
            assert bytes(b["seqs"][b["off"][i]:b["off"][i + 1]]) == seq
            if b["quals"] is not None:
                assert bytes(b["quals"][b["off"][i]:b["off"][i + 1]]) == (qual or b"")
            else:
                assert not qual
            if b["comments"] is not None:
                cf, co = b["comments"]
                assert bytes(cf[co[i]:co[i + 1]]) == comment
            else:
                assert comment == b""
    assert reader.read(CHUNK) is None


CHUNK = 30000


def _fastq_text(n, seed, suffix=b"", crlf=False, with_comments=True, lens=(30, 260)):
    rng = np.random.default_rng(seed)
    nl = b"\r\n" if crlf else b"\n"
    out = b""
    for i in range(n):
        l = int(rng.integers(*lens))
        seq = bytes(rng.choice(np.frombuffer(b"ACGTNacgt", dtype=np.uint8), size=l))
        qual = bytes(rng.integers(33, 74, size=l, dtype=np.uint8))     # '@' (64) and '+' (43) occur at line starts
        c = (b" BC:Z:%d extra words" % i) if with_comments and i % 3 else b""
        out += b"@read%d%s%s" % (i, suffix, c) + nl + seq + nl + b"+" + nl + qual + nl
    return out


@pytest.mark.parametrize("threads", [1, 7])
def test_fastq_plain_gzip_bgzf(tmp_path, threads):
    lib = cm.backend_lib("emu")
    text = _fastq_text(3001, 5, suffix=b"/1")
    exp = _py_batches([_py_records(text)], CHUNK)
    assert len(exp) > 5 and all(len(b) % 2 == 0 for b in exp[:-1])
    for name, data in (("a.fq", text), ("a.fq.gz", gzip.compress(text)), ("a.bgzf.gz", _bgzf(text))):
        p = tmp_path / name
        p.write_bytes(data)
        r = L.FastqReader(str(p), n_threads=threads, lib=lib)
        _check(r, exp)
        r.close()


def test_fastq_two_files_fasta_crlf(tmp_path):
    lib = cm.backend_lib("emu")
    t1, t2 = _fastq_text(1500, 7, suffix=b"/1", with_comments=False), _fastq_text(1500, 8, suffix=b"/2", crlf=True, with_comments=False)
    (tmp_path / "m1.fq").write_bytes(t1); (tmp_path / "m2.fq.gz").write_bytes(gzip.compress(t2))
    r = L.FastqReader(str(tmp_path / "m1.fq"), str(tmp_path / "m2.fq.gz"), n_threads=4, lib=lib)
    _check(r, _py_batches([_py_records(t1), _py_records(t2)], CHUNK))
    r.close()
    rng = np.random.default_rng(3)                             # multi-line FASTA, no qualities, a long record
    fa = b""
    for i in range(40):
        s = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=int(rng.integers(50, 9000))))
        fa += b">ctg%d some description\n" % i + b"\n".join(s[k:k + 60] for k in range(0, len(s), 60)) + b"\n"
    (tmp_path / "x.fa").write_bytes(fa)
    r = L.FastqReader(str(tmp_path / "x.fa"), n_threads=3, lib=lib)
    _check(r, _py_batches([_py_records(fa)], CHUNK))
    r.close()
    with pytest.raises(L.B200Error):
        L.FastqReader(str(tmp_path / "missing.fq"), lib=lib)


@pytest.mark.gpu
def test_fastq_files_to_sam_matches_reference_program(tmp_path):
    """FASTQ files -> SAM through the reader and b200_reads_to_sam, batch by batch, against `bwa_cpu mem` on the same files."""
    import test_cli
    if not test_cli.have_cli:
        pytest.skip("cli/_build not built")
    contigs = synth.make_genome([300_000, 120_000], seed=78)
    fa = str(tmp_path / "ref.fa")
    synth.write_fasta(fa, contigs)
    test_cli._run(test_cli.BWA_CPU, ["index", fa])
    pairs, _ = synth.simulate_pairs(contigs, 4000, seed=8, damaged_frac=0.1, damage="periodic")
    o = np.arange(4001, dtype=np.int64) * 150
    rng = np.random.default_rng(2)
    q = rng.integers(35, 74, size=4000 * 150, dtype=np.uint8)
    synth.write_fastq(str(tmp_path / "p1.fq"), np.ascontiguousarray(pairs[0::2].reshape(-1)), o, [f"p{i}/1" for i in range(4000)], q)
    synth.write_fastq(str(tmp_path / "p2.fq"), np.ascontiguousarray(pairs[1::2].reshape(-1)), o, [f"p{i}/2" for i in range(4000)], q)
    chunk = 500_000
    exp = test_cli._run(test_cli.BWA_CPU, ["mem", "-t", "4", "-K", str(chunk), fa, str(tmp_path / "p1.fq"), str(tmp_path / "p2.fq")])
    exp_body = b"\n".join(l for l in exp.split(b"\n") if not l.startswith(b"@"))
    lib = cm.backend_lib("cuda")
    ctx = L.Context(0, lib)
    h = L.HostIndex(contigs, lib)
    ctx.upload_host_index(h)
    opt = abi.default_opt(); opt.flag |= abi.MEM_F_PE; opt.n_threads = 4
    r = L.FastqReader(str(tmp_path / "p1.fq"), str(tmp_path / "p2.fq"), n_threads=4, lib=lib)
    got, n_processed = b"", 0
    while True:
        b = r.read(chunk)
        if b is None:
            break
        nf, no = b["names"]
        sam, _, _ = L.reads_to_sam([ctx], opt, h.names, b["seqs"], b["off"], nf, no, quals=b["quals"], n_processed=n_processed)
        got += sam.tobytes(); n_processed += b["n"]
    r.close()
    assert got == exp_body
    # the same as one streaming call: reader thread | devices | writer thread
    out = str(tmp_path / "out.sam")
    n, st = L.fastq_to_sam([ctx], opt, h.names, str(tmp_path / "p1.fq"), str(tmp_path / "p2.fq"), chunk_bases=chunk, out_path=out, n_threads=4)
    assert n == 8000 and open(out, "rb").read() == exp_body and st[4] >= 2
    ctx.close()
