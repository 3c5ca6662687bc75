"""Synthetic genomes and reads for the BASELINE.json configs.

No wgsim / samtools exist on the box, so the workload generator lives here.  Everything is seeded
and numpy-vectorised so that the 10 M-read config is produced in seconds.

Encoding: nt4 codes 0..3 = A,C,G,T (4 = N) as uint8 numpy arrays; `to_ascii` converts for FASTA/FASTQ.
"""
from __future__ import annotations

import numpy as np

GENOME_SEED = 20261104
_ASCII = np.frombuffer(b"ACGTN", dtype=np.uint8)


def to_ascii(nt4: np.ndarray) -> np.ndarray:
    return _ASCII[nt4]


def make_genome(contig_lens, seed: int = GENOME_SEED, repeat_frac: float = 0.0):
    """i.i.d. uniform A/C/G/T contigs.  repeat_frac > 0 overwrites that fraction of the genome with
    copies of 50 repeat families (300-6000 bp, 0-5 % divergence): the repeat-stress variant."""
    rng = np.random.default_rng(seed)
    contigs = [rng.integers(0, 4, size=int(n), dtype=np.uint8) for n in contig_lens]
    if repeat_frac > 0:
        total = sum(len(c) for c in contigs)
        budget = int(total * repeat_frac)
        fams = []
        for _ in range(50):
            fl = int(rng.integers(300, 6001))
            fams.append((rng.integers(0, 4, size=fl, dtype=np.uint8), float(rng.uniform(0, 0.05))))
        while budget > 0:
            fam, div = fams[int(rng.integers(0, len(fams)))]
            c = contigs[int(rng.integers(0, len(contigs)))]
            if len(c) <= len(fam) + 1:
                continue
            pos = int(rng.integers(0, len(c) - len(fam)))
            cp = fam.copy()
            nm = rng.random(len(cp)) < div
            cp[nm] = (cp[nm] + rng.integers(1, 4, size=int(nm.sum()))) % 4
            if rng.random() < 0.5:
                cp = (3 - cp)[::-1]
            c[pos:pos + len(cp)] = cp
            budget -= len(cp)
    return contigs


def write_fasta(path: str, contigs, names=None, width: int = 60) -> None:
    with open(path, "wb") as f:
        for i, c in enumerate(contigs):
            nm = names[i] if names else f"chr{i + 1}"
            f.write(f">{nm}\n".encode())
            a = to_ascii(c)
            n_full = len(a) // width
            if n_full:
                body = np.empty((n_full, width + 1), dtype=np.uint8)
                body[:, :width] = a[: n_full * width].reshape(n_full, width)
                body[:, width] = 10
                f.write(body.tobytes())
            if len(a) % width:
                f.write(a[n_full * width:].tobytes() + b"\n")


def simulate_reads(contigs, n_reads: int, read_len: int = 150, sub_rate: float = 0.01,
                   indel_rate: float = 0.00015, seed: int = 1001, n_rate: float = 0.0):
    """Fixed-length single-end reads: uniform start over the concatenated contigs (reads that would
    span a contig end are re-drawn inside the contig), strand 50/50, per-base substitutions at
    sub_rate, 1-bp indels at indel_rate.  Returns (reads[n, read_len] nt4 uint8, truth dict)."""
    rng = np.random.default_rng(seed)
    lens = np.array([len(c) for c in contigs], dtype=np.int64)
    cat = np.concatenate(contigs)
    starts_c = np.concatenate([[0], np.cumsum(lens)])[:-1]
    cid = rng.choice(len(contigs), size=n_reads, p=lens / lens.sum())
    pad = 8
    pos = (rng.random(n_reads) * (lens[cid] - read_len - pad)).astype(np.int64)
    gpos = starts_c[cid] + pos
    reads = np.empty((n_reads, read_len), dtype=np.uint8)
    CH = 1 << 20
    ar = np.arange(read_len, dtype=np.int64)
    for b in range(0, n_reads, CH):
        e = min(n_reads, b + CH)
        reads[b:e] = cat[gpos[b:e, None] + ar[None, :]]
    # indels on a sparse subset of reads (python loop over the affected reads only)
    if indel_rate > 0:
        n_ev = rng.binomial(n_reads * read_len, indel_rate)
        ev_read = rng.integers(0, n_reads, size=n_ev)
        ev_pos = rng.integers(1, read_len - 1, size=n_ev)
        ev_ins = rng.random(n_ev) < 0.5
        ev_base = rng.integers(0, 4, size=n_ev, dtype=np.uint8)
        for r, p, ins, bb in zip(ev_read, ev_pos, ev_ins, ev_base):
            src = cat[gpos[r]: gpos[r] + read_len + pad]
            cur = reads[r]
            if ins:
                reads[r] = np.concatenate([cur[:p], [bb], cur[p:-1]])
            else:
                reads[r] = np.concatenate([cur[:p], cur[p + 1:], src[read_len:read_len + 1]])
    # substitutions (sparse positions)
    total = n_reads * read_len
    n_sub = rng.binomial(total, sub_rate)
    flat = reads.reshape(-1)
    sp = rng.integers(0, total, size=n_sub)
    flat[sp] = (flat[sp] + rng.integers(1, 4, size=n_sub).astype(np.uint8)) % 4
    if n_rate > 0:
        n_n = rng.binomial(total, n_rate)
        flat[rng.integers(0, total, size=n_n)] = 4
    # strand
    rev = rng.random(n_reads) < 0.5
    idx = np.nonzero(rev)[0]
    for b in range(0, len(idx), CH):
        ii = idx[b:b + CH]
        blk = reads[ii][:, ::-1]
        reads[ii] = np.where(blk < 4, 3 - blk, blk)
    return reads, {"contig": cid, "pos": pos, "rev": rev}


def simulate_long_reads(contigs, n_reads: int, min_len: int = 1000, max_len: int = 10000,
                        sub: float = 0.03, ins: float = 0.02, dele: float = 0.02, seed: int = 1005,
                        junk_frac: float = 0.10):
    """Config-5 style long reads (ragged): returns (flat nt4 uint8, offsets int64[n+1])."""
    rng = np.random.default_rng(seed)
    lens = np.array([len(c) for c in contigs], dtype=np.int64)
    out, off = [], [0]
    for _ in range(n_reads):
        L = int(rng.integers(min_len, max_len + 1))
        c = int(rng.choice(len(contigs), p=lens / lens.sum()))
        L = min(L, int(lens[c]) - 16)
        p = int(rng.integers(0, lens[c] - L))
        s = contigs[c][p:p + L].copy()
        m = rng.random(L) < sub
        s[m] = (s[m] + rng.integers(1, 4, size=int(m.sum()))) % 4
        keep = rng.random(L) >= dele
        s = s[keep]
        ip = np.nonzero(rng.random(len(s)) < ins)[0]
        if len(ip):
            s = np.insert(s, ip, rng.integers(0, 4, size=len(ip), dtype=np.uint8))
        if rng.random() < junk_frac:
            s = np.concatenate([s, rng.integers(0, 4, size=int(rng.integers(200, 501)), dtype=np.uint8)])
        if rng.random() < 0.5:
            s = (3 - s)[::-1]
        out.append(s.astype(np.uint8))
        off.append(off[-1] + len(s))
    return np.concatenate(out), np.array(off, dtype=np.int64)


def simulate_pairs(contigs, n_pairs: int, read_len: int = 150, ins_mean: float = 500.0, ins_sd: float = 50.0,
                   sub_rate: float = 0.01, damaged_frac: float = 0.05, seed: int = 1003, damage: str = "window"):
    """Config-3 style FR pairs, interleaved (2*n_pairs, read_len); a damaged_frac of pairs have one damaged mate.
    damage="window": 30 % substitutions over a random 100-bp window (such a mate usually still
    seeds from its clean part, so few rescues run).  damage="periodic": one substitution every 14 bases, i.e. no exact
    match of min_seed_len = 19 survives: the mate gets no region at all and can only be placed by mem_matesw."""
    rng = np.random.default_rng(seed)
    lens = np.array([len(c) for c in contigs], dtype=np.int64)
    cat = np.concatenate(contigs)
    starts_c = np.concatenate([[0], np.cumsum(lens)])[:-1]
    cid = rng.choice(len(contigs), size=n_pairs, p=lens / lens.sum())
    isz = np.clip(rng.normal(ins_mean, ins_sd, size=n_pairs).astype(np.int64), read_len + 10, None)
    pos = (rng.random(n_pairs) * (lens[cid] - isz - 8)).astype(np.int64)
    g = starts_c[cid] + pos
    ar = np.arange(read_len, dtype=np.int64)
    r1 = cat[g[:, None] + ar[None, :]]
    r2f = cat[(g + isz - read_len)[:, None] + ar[None, :]]
    r2 = 3 - r2f[:, ::-1]
    flip = rng.random(n_pairs) < 0.5            # fragment from the reverse strand
    a = np.where(flip[:, None], r2, r1)
    b = np.where(flip[:, None], r1, r2)
    reads = np.empty((2 * n_pairs, read_len), dtype=np.uint8)
    reads[0::2] = a
    reads[1::2] = b
    flat = reads.reshape(-1)
    total = flat.size
    n_sub = rng.binomial(total, sub_rate)
    sp = rng.integers(0, total, size=n_sub)
    flat[sp] = (flat[sp] + rng.integers(1, 4, size=n_sub).astype(np.uint8)) % 4
    dmg = np.nonzero(rng.random(n_pairs) < damaged_frac)[0]
    for p in dmg:
        r = 2 * p + int(rng.integers(0, 2))
        if damage == "periodic":
            ph = int(rng.integers(0, 14))
            reads[r, ph::14] = (reads[r, ph::14] + rng.integers(1, 4, size=len(reads[r, ph::14]))) % 4
            continue
        w0 = int(rng.integers(0, read_len - 100 + 1))
        m = rng.random(100) < 0.30
        seg = reads[r, w0:w0 + 100]
        seg[m] = (seg[m] + rng.integers(1, 4, size=int(m.sum()))) % 4
    return reads, {"contig": cid, "pos": pos, "isize": isz, "flip": flip}


# ---- counter-based read simulator: read i depends on (seed, i) only, so any slice of a 10^8-read set can be produced
# on its own -- by numpy on the host (the CPU arm's sample) or by torch on the GPU (the full set in seconds) -- with
# identical results: everything is wrap-around int64 arithmetic, which both libraries implement the same way.
_M1, _M2, _G = 0xBF58476D1CE4E5B9 - (1 << 64), 0x94D049BB133111EB - (1 << 64), 0x9E3779B97F4A7C15 - (1 << 64)


def _lsr(x, s):                                   # logical shift right of int64
    return (x >> s) & ((1 << (64 - s)) - 1)


def _mix(x):                                      # splitmix64 finaliser on int64 arrays / tensors
    z = x + _G
    z = (z ^ _lsr(z, 30)) * _M1
    z = (z ^ _lsr(z, 27)) * _M2
    return z ^ _lsr(z, 31)


def counter_reads(xp, genome, n_contigs: int, contig_len: int, i0: int, i1: int, read_len: int = 150, seed: int = 1004,
                  sub_per_64k: int = 655, indel_per_64k_read: int = 1475):
    """Reads i0 .. i1 of the fixed read set `seed` over a genome of n_contigs equal contigs (concatenated nt4 array /
    tensor `genome`): uniform start inside a contig, strand 50/50, substitutions at sub_per_64k / 65536 per base (1 %),
    one 1-bp insertion or deletion in indel_per_64k_read / 65536 of the reads (2.25 % = 0.015 % per base).
    xp = numpy or torch; `genome` on the device the reads should be made on.  Returns (reads [n, read_len] uint8,
    contig, pos, rev)."""
    is_t = xp.__name__ == "torch"
    ar = (lambda a, b: xp.arange(a, b, dtype=xp.int64, device=genome.device)) if is_t else (lambda a, b: xp.arange(a, b, dtype=xp.int64))
    i = ar(i0, i1)
    h = _mix(i * 2 + seed * 1000003)
    h63 = h & ((1 << 63) - 1)
    contig = h63 % n_contigs
    pos = (h63 // n_contigs) % (contig_len - read_len - 8)
    rev = _lsr(h, 63)
    g = contig * contig_len + pos
    h2 = _mix(i * 2 + 1 + seed * 1000003)
    has_indel = (h2 & 0xffff) < indel_per_64k_read
    is_ins = (_lsr(h2, 16) & 1) == 1
    ip = 1 + (_lsr(h2, 17) & 0xffff) % (read_len - 2)                # event position 1 .. read_len - 2
    ibase = _lsr(h2, 40) & 3
    j = ar(0, read_len)[None, :]
    # source offset of read base j: deletion skips source base ip, insertion repeats from ip on
    dele = (has_indel & ~is_ins)[:, None]
    inse = (has_indel & is_ins)[:, None]
    shift = (dele & (j >= ip[:, None])) * 1 - (inse & (j > ip[:, None])) * 1
    src = g[:, None] + j + shift
    reads = genome[src] if not is_t else genome[src.reshape(-1)].reshape(src.shape)
    reads = reads.to(xp.int64) if is_t else reads.astype(xp.int64)
    at_ins = inse & (j == ip[:, None])
    reads = xp.where(at_ins, ibase[:, None] + 0 * j, reads)
    hb = _mix((i[:, None] * read_len + j) * 4 + 3 + seed * 7000003)
    sub = (hb & 0xffff) < sub_per_64k
    reads = xp.where(sub, (reads + 1 + (_lsr(hb, 16) & 0xffff) % 3) % 4, reads)
    r2 = 3 - (xp.flip(reads, (1,)) if is_t else reads[:, ::-1])
    reads = xp.where((rev == 1)[:, None], r2, reads)
    reads = reads.to(xp.uint8) if is_t else reads.astype(xp.uint8)
    return reads, contig, pos, rev


def write_fastq(path: str, reads_nt4, off, names, quals=None, comments=None) -> None:
    """FASTQ file of reads given as flat nt4 codes + offsets; quals: flat ASCII bytes with the same offsets (default 'I')."""
    asc = to_ascii(np.asarray(reads_nt4))
    with open(path, "wb") as f:
        for i, nm in enumerate(names):
            s = asc[off[i]:off[i + 1]].tobytes()
            q = quals[off[i]:off[i + 1]].tobytes() if quals is not None else b"I" * len(s)
            c = (" " + comments[i]) if comments is not None and comments[i] else ""
            f.write(b"@" + nm.encode() + c.encode() + b"\n" + s + b"\n+\n" + q + b"\n")
