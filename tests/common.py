"""Shared helpers for the parity tests.

Backends:
  "emu"  -- tests/emu/_build/libbwamem_b200_emu.so: the product's kernels compiled against the SIMT emulator
            (logic check in the GPU-less container; `-m "not gpu"`).
  "cuda" -- bwa-mem-gpu_b200/libbwamem_b200.so on a real GPU (`-m gpu`).
The checker is the unmodified reference CPU code (oracle/_ref/libbwaref.so) when present, and the committed
golden vectors under tests/golden/ (generated from it by tests/golden/make_golden.py) otherwise.
"""
from __future__ import annotations

import functools
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.path.join(ROOT, "tests", "_data")
GOLDEN = os.path.join(ROOT, "tests", "golden")
EMU_SO = os.environ.get("B200_EMU_SO") or os.path.join(ROOT, "tests", "emu", "_build", "libbwamem_b200_emu.so")      # B200_EMU_SO: e.g. the AddressSanitizer build of scripts/emu_asan.sh

from bwa_mem_gpu_b200 import abi, lib as L, synth  # noqa: E402
import refapi  # noqa: E402


def have_ref():
    return refapi.available()


@functools.lru_cache(maxsize=None)
def reflib():
    return refapi.RefLib()


@functools.lru_cache(maxsize=None)
def backend_lib(backend: str):
    if backend == "emu":
        if not os.path.exists(EMU_SO):
            subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "bwa-mem-gpu_b200", "csrc"), "emu", "-j8"])
        return L.load(EMU_SO)
    return L.load()


GENOMES = {
    # name: (contig lengths, seed, repeat_frac)
    "tiny": ([20000, 12000, 7000], 5, 0.0),
    "small": ([200000, 100000], 7, 0.0),
    "rep": ([150000, 90000], 9, 0.08),
    # explicit duplications (secondary hits, sub_n, XA within and beyond max_XA_hits) and an ALT contig (chr3) that is a
    # diverged copy of a chr1 segment: see dup_genome()
    "dup": ([60000, 40000, 5000], 11, 0.0),
}
ALT_CONTIGS = {"dup": [2]}


def dup_genome():
    lens, seed, _ = GENOMES["dup"]
    rng = np.random.default_rng(seed)
    c = [rng.integers(0, 4, size=int(n), dtype=np.uint8) for n in lens]

    def mutated(x, div):
        y = x.copy()
        m = rng.random(len(y)) < div
        y[m] = (y[m] + rng.integers(1, 4, size=int(m.sum()))) % 4
        return y
    a = c[0][1000:3000].copy()                               # 2 kb, 4 more copies at 0 - 2 % divergence
    for k, (ci, pos, div) in enumerate([(0, 20000, 0.0), (1, 5000, 0.005), (1, 15000, 0.01), (0, 40000, 0.02)]):
        y = mutated(a, div)
        c[ci][pos:pos + len(y)] = (3 - y)[::-1] if k & 1 else y
    b = c[1][30000:30400].copy()                             # 400 bp, 8 copies: more than max_XA_hits
    for k in range(7):
        c[0][45000 + 1000 * k:45400 + 1000 * k] = mutated(b, 0.003)
    c[2][:] = mutated(c[0][10000:15000], 0.01)               # the ALT contig
    return c


@functools.lru_cache(maxsize=None)
def genome(name: str):
    if name == "dup":
        return dup_genome()
    lens, seed, rep = GENOMES[name]
    return synth.make_genome(lens, seed=seed, repeat_frac=rep)


@functools.lru_cache(maxsize=None)
def ref_index(name: str):
    os.makedirs(DATA, exist_ok=True)
    pre = os.path.join(DATA, f"ref_{name}")
    if not os.path.exists(pre + ".sa"):
        synth.write_fasta(pre + ".fa", genome(name))
        reflib().index_build(pre + ".fa", pre)
    if name in ALT_CONTIGS and not os.path.exists(pre + ".alt"):     # bns_restore reads <prefix>.alt (bntseq.c:168-195)
        with open(pre + ".alt", "w") as f:
            f.write("".join(f"chr{i + 1}\n" for i in ALT_CONTIGS[name]))
    return reflib().index_load(pre)


@functools.lru_cache(maxsize=None)
def context(backend: str, name: str):
    """Device context holding the index of genome `name`, built by OUR host index builder."""
    lib = backend_lib(backend)
    h = L.HostIndex(genome(name), lib)
    for i in ALT_CONTIGS.get(name, []):
        h.ann_alt[i] = 1
    ctx = L.Context(0, lib)
    ctx.upload_host_index(h)
    ctx._host_index = h
    return ctx


def reads_for(name: str, n: int, seed: int = 3, read_len: int = 150, sub=0.01, indel=0.002, n_rate=0.0):
    r, _ = synth.simulate_reads(genome(name), n, read_len=read_len, sub_rate=sub, indel_rate=indel, seed=seed, n_rate=n_rate)
    seqs = np.ascontiguousarray(r.reshape(-1))
    off = np.arange(0, n + 1, dtype=np.int64) * read_len
    return seqs, off


def ragged_reads(name: str, seed: int = 17):
    """Edge cases: empty read, shorter than min_seed_len, all-N, N-interrupted, ASCII mixed case, long."""
    rng = np.random.default_rng(seed)
    g = np.concatenate(genome(name))
    parts = []
    parts.append(np.zeros(0, dtype=np.uint8))                       # empty
    parts.append(g[100:110].copy())                                 # < min_seed_len
    parts.append(np.full(40, 4, dtype=np.uint8))                    # all N
    a = g[500:700].copy(); a[60] = 4; a[61] = 4; a[150] = 4
    parts.append(a)                                                 # N inside
    parts.append(g[1000:1019].copy())                               # exactly min_seed_len
    parts.append(g[2000:2400].copy())                               # 400 bp exact
    b = g[3000:3600].copy()
    m = rng.random(len(b)) < 0.03
    b[m] = (b[m] + 1) % 4
    parts.append(b)                                                 # 600 bp, 3 % subs (just below the seed-SW threshold, bwamem.c:609)
    c = (3 - g[5000:5250])[::-1].copy()
    parts.append(c)                                                 # reverse strand
    parts.append(rng.integers(0, 4, size=200, dtype=np.uint8))      # unrelated random
    d = np.concatenate([g[6000:6100], g[9000:9100]])
    parts.append(d)                                                 # chimeric
    seqs = np.concatenate(parts).astype(np.uint8)
    off = np.concatenate([[0], np.cumsum([len(p) for p in parts])]).astype(np.int64)
    return np.ascontiguousarray(seqs), off


def golden_path(name):
    return os.path.join(GOLDEN, name)


KSW_XBYTE, KSW_XSTOP, KSW_XSUBO, KSW_XSTART = 0x10000, 0x20000, 0x40000, 0x80000


def sw_problems(rng, n, maxq=160, maxt=420, sat=False):
    """Local-alignment problems in the shape mem_matesw / mem_seed_sw produce (bwamem_pair.c:150-162, bwamem.c:640-672):
    a target window holding zero, one or two diverged copies of (a piece of) the query."""
    qs, ts = [], []
    for _ in range(n):
        lq = int(rng.integers(1, maxq))
        q = rng.integers(0, 4, size=lq, dtype=np.uint8)
        if rng.random() < 0.1 and lq > 3:
            q[int(rng.integers(0, lq))] = 4
        parts = [rng.integers(0, 4, size=int(rng.integers(0, maxt // 3)), dtype=np.uint8)]
        for _c in range(int(rng.integers(0, 3))):
            a, b = sorted(rng.integers(0, lq + 1, size=2))
            piece = []
            for x in q[a:b]:
                r = rng.random()
                if r < 0.05: piece.append(int(rng.integers(0, 4)))
                elif r < 0.065: continue
                elif r < 0.08: piece.extend([int(x) & 3, int(rng.integers(0, 4))])
                else: piece.append(int(x) & 3 if x < 4 else 4)
            parts.append(np.array(piece, dtype=np.uint8))
            parts.append(rng.integers(0, 4, size=int(rng.integers(0, maxt // 4)), dtype=np.uint8))
        t = np.concatenate(parts)[:maxt]
        if sat and rng.random() < 0.3:                      # identical long pair: byte scores saturate at 255
            q = rng.integers(0, 4, size=int(rng.integers(250, 300)), dtype=np.uint8); t = q.copy()
        qs.append(q); ts.append(t)
    ql = np.array([len(x) for x in qs], np.int32); tl = np.array([len(x) for x in ts], np.int32)
    qoff = np.concatenate([[0], np.cumsum(ql)[:-1]]).astype(np.int64)
    toff = np.concatenate([[0], np.cumsum(tl)[:-1]]).astype(np.int64)
    q = np.ascontiguousarray(np.concatenate(qs + [np.zeros(1, np.uint8)]).astype(np.uint8))
    t = np.ascontiguousarray(np.concatenate(ts + [np.zeros(1, np.uint8)]).astype(np.uint8))
    return ql, qoff, q, tl, toff, t


def reflib_pestat(idx, opt, regs, cnt):
    return idx.pestat(opt, regs, cnt)
