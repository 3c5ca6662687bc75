"""One process per GPU: the read batch is cut into contiguous blocks (whole pairs), one per rank, the FM-index is
replicated, and no collective touches the data path.  The only exchanges are host-side and per batch:

  * results are gathered on rank 0 in input order (what the reference's single process hands to worker2 /
    SAM output, bwamem.c:1240);
  * paired-end: mem_pestat (bwamem_pair.c:72) is a statistic over ALL pairs of the batch, so the region lists are
    gathered on rank 0, the host program's mem_pestat runs there once, and the 4 x mem_pestat_t (128 bytes) are
    broadcast before every rank runs the mate rescue of its own block.

torch.distributed is plumbing only (gloo for host arrays; the device work is the C ABI of libbwamem_b200).
"""
from __future__ import annotations

import numpy as np

from .abi import ALNREG_DTYPE, PESTAT_DTYPE, MEM_F_PE


def block_bounds(n_reads: int, world: int, pair_aligned: bool = False):
    """[lo, hi) of every rank's block; with pair_aligned the cuts fall on even read indices."""
    b = [n_reads * r // world for r in range(world + 1)]
    if pair_aligned:
        b = [x & ~1 for x in b[:-1]] + [n_reads]
    return [(b[r], b[r + 1]) for r in range(world)]


def _gather_arrays(arr: np.ndarray, dist, rank: int, world: int):
    """Variable-length gather of a 1-D numpy array to rank 0 (list of per-rank arrays there, None elsewhere)."""
    import torch
    raw = torch.from_numpy(np.ascontiguousarray(arr).view(np.uint8).reshape(-1).copy())
    sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([raw.numel()], dtype=torch.int64))
    mx = int(max(int(s) for s in sizes))
    pad = torch.zeros(mx, dtype=torch.uint8)
    pad[: raw.numel()] = raw
    bufs = [torch.zeros(mx, dtype=torch.uint8) for _ in range(world)] if rank == 0 else None
    dist.gather(pad, bufs, dst=0)
    if rank != 0:
        return None
    return [bufs[r][: int(sizes[r])].numpy().view(arr.dtype) for r in range(world)]


def align_sharded(ctx, opt, seqs: np.ndarray, off: np.ndarray, dist=None, pestat_fn=None):
    """mem_align1_core for every read of the batch (+ the mate-rescue stage when opt.flag has MEM_F_PE), this rank
    computing its own block.  Returns (regs, counts) for the WHOLE batch on rank 0, (None, None) elsewhere.

    dist      : an initialised torch.distributed module (any backend able to move CPU tensors), or None = 1 rank
    pestat_fn : callable(regs, counts) -> 4 x PESTAT_DTYPE, the host program's mem_pestat; needed on rank 0 for MEM_F_PE
    """
    rank = dist.get_rank() if dist is not None else 0
    world = dist.get_world_size() if dist is not None else 1
    n = len(off) - 1
    pe = bool(opt.flag & MEM_F_PE)
    if pe and not (opt.flag & 0x20):                    # rank 0 needs pestat_fn; every rank learns of a missing one BEFORE any gather
        missing = int(rank == 0 and pestat_fn is None)
        if world > 1:
            import torch
            t = torch.tensor([missing], dtype=torch.int32)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            missing = int(t[0])
        if missing:
            raise ValueError("paired-end batches need the host program's mem_pestat (pestat_fn) on rank 0")
    lo, hi = block_bounds(n, world, pe)[rank]
    my_off = np.ascontiguousarray(off[lo: hi + 1])
    ctx.batch_upload(seqs, my_off)
    ctx.batch_run(opt)
    cap = 2 * (hi - lo) + 64

    def download():
        nonlocal cap
        while True:
            regs, cnt = ctx.batch_download(hi - lo, cap)
            if regs is not None:
                return regs, cnt
            cap = cnt                                   # -(needed) came back as the second value

    def gather(regs, cnt):
        if world == 1:
            return regs, cnt
        rr = _gather_arrays(regs, dist, rank, world)
        cc = _gather_arrays(cnt, dist, rank, world)
        if rank != 0:
            return None, None
        return np.concatenate(rr).view(ALNREG_DTYPE), np.concatenate(cc).view(np.int32)

    regs, cnt = download()
    if not pe or (opt.flag & 0x20):                     # MEM_F_NO_RESCUE
        return gather(regs, cnt)
    all_regs, all_cnt = gather(regs, cnt)
    pes = np.zeros(4, dtype=PESTAT_DTYPE)
    if rank == 0:
        pes[:] = pestat_fn(all_regs, all_cnt)
    if world > 1:
        import torch
        t = torch.from_numpy(pes.view(np.uint8).copy())
        dist.broadcast(t, src=0)
        pes = t.numpy().view(PESTAT_DTYPE).copy()
    ctx.batch_rescue(opt, pes)
    regs, cnt = download()
    return gather(regs, cnt)
