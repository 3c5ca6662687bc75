"""Workload side of bench.py: the synthetic reference of a BASELINE.json config and its `bwa index` arrays, kept as a
cache directory of plain .npy files so that both bench arms -- ours and the reference's CPU code -- run on the same index
without the reference arm ever loading the product library.

`bwa index` on 3.1 Gbp takes hours of host time; its arrays (bwt_t.bwt with Occ checkpoints, bwtindex.c:149-172; the
sampled SA, bwt.c:62-84; the 2-bit .pac, bntseq.c:230) are determined by the text, so they are computed by the device
builder (b200_index_build_device, byte-identical to `bwa index`: tests/test_index.py) in a few seconds and written out.

    python -m bwa_mem_gpu_b200.bench_index --contigs 24 --contig-len 129166667 --out DIR [--device 0]

The genome is i.i.d. uniform A/C/G/T from a counter-based 64-bit generator (splitmix64 finaliser of the base's index,
seed 20261104), so any part of it can be re-derived anywhere.
"""
from __future__ import annotations

import argparse
import json
import os
import time

import numpy as np

GENOME_SEED = 20261104
FILES = ("pac.npy", "bwt.npy", "sa.npy", "meta.json")


def cache_dir(root: str, contigs: int, contig_len: int) -> str:
    return os.path.join(root, "tests", "_data", f"bench_index_{contigs}x{contig_len}")


def have(d: str) -> bool:
    return all(os.path.exists(os.path.join(d, f)) for f in FILES)


def load(d: str, mmap: bool = True):
    meta = json.load(open(os.path.join(d, "meta.json")))
    mode = "r" if mmap else None
    return meta, np.load(os.path.join(d, "pac.npy"), mmap_mode=mode), np.load(os.path.join(d, "bwt.npy"), mmap_mode=mode), \
        np.load(os.path.join(d, "sa.npy"), mmap_mode=mode)


def genome_pac_torch(contigs: int, contig_len: int, device):
    """2-bit packed forward reference (bntseq.c:230: base i in byte i >> 2 at bits (~i & 3) * 2) made on the GPU."""
    import torch
    from .synth import _mix
    l_pac = contigs * contig_len
    pac = torch.zeros(l_pac // 4 + 1, dtype=torch.uint8, device=device)
    CH = 1 << 28
    for b in range(0, l_pac, CH):
        e = min(l_pac, b + CH)
        e4 = e if e == l_pac else e                           # CH is a multiple of 4
        k = torch.arange(b, e4, dtype=torch.int64, device=device)
        base = _mix(k + GENOME_SEED * 7919) & 3
        pad = (-len(base)) % 4
        if pad:
            base = torch.cat([base, torch.zeros(pad, dtype=torch.int64, device=device)])
        q = base.view(-1, 4)
        pac[b // 4: b // 4 + q.shape[0]] = ((q[:, 0] << 6) | (q[:, 1] << 4) | (q[:, 2] << 2) | q[:, 3]).to(torch.uint8)
        del k, base, q
    return pac, l_pac


def unpack_pac_torch(pac_t, l_pac: int):
    """nt4 codes of the whole forward reference as a uint8 tensor on pac_t's device."""
    import torch
    sh = torch.tensor([6, 4, 2, 0], dtype=torch.uint8, device=pac_t.device)
    g = ((pac_t[:, None] >> sh[None, :]) & 3).reshape(-1)
    return g[:l_pac].contiguous()


def unpack_pac_numpy(pac, lo: int, hi: int):
    """nt4 codes of bases lo .. hi of the forward reference."""
    b0, b1 = lo // 4, (hi + 3) // 4
    p = np.asarray(pac[b0:b1])
    g = np.empty(len(p) * 4, dtype=np.uint8)
    g[0::4] = p >> 6; g[1::4] = (p >> 4) & 3; g[2::4] = (p >> 2) & 3; g[3::4] = p & 3
    return g[lo - 4 * b0: hi - 4 * b0]


def build(contigs: int, contig_len: int, out: str, device: int = 0):
    import torch
    from . import lib as L
    t0 = time.time()
    dev = torch.device("cuda", device)
    pac_t, l_pac = genome_pac_torch(contigs, contig_len, dev)
    pac = pac_t.cpu().numpy()
    del pac_t
    torch.cuda.empty_cache()
    t1 = time.time()
    ctx = L.Context(device)
    hidx = ctx.build_index_device_pac(pac, [contig_len] * contigs)
    t2 = time.time()
    os.makedirs(out, exist_ok=True)
    np.save(os.path.join(out, "pac.npy"), pac)
    np.save(os.path.join(out, "bwt.npy"), hidx.bwt_words())
    np.save(os.path.join(out, "sa.npy"), hidx.sa())
    meta = {"contigs": contigs, "contig_len": contig_len, "l_pac": l_pac, "primary": hidx.primary, "L2": hidx.L2,
            "seq_len": hidx.seq_len, "bwt_size": hidx.bwt_size, "n_sa": hidx.n_sa, "sa_intv": hidx.sa_intv,
            "genome": f"i.i.d. uniform ACGT, splitmix64 of the base index, seed {GENOME_SEED}",
            "genome_s": t1 - t0, "index_build_s": t2 - t1, "write_s": None,
            "builder": "b200_index_build_device (byte-identical to bwa index: tests/test_index.py)"}
    meta["write_s"] = time.time() - t2
    json.dump(meta, open(os.path.join(out, "meta.json"), "w"))
    ctx.close()
    return meta


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--contigs", type=int, required=True)
    ap.add_argument("--contig-len", type=int, required=True)
    ap.add_argument("--out", required=True)
    ap.add_argument("--device", type=int, default=0)
    a = ap.parse_args()
    print(json.dumps(build(a.contigs, a.contig_len, a.out, a.device)))
