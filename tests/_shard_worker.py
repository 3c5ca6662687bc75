"""Worker of tests/test_shard.py: one rank of a world_size-N job (gloo on CPU with the emulated kernels, or one GPU
per rank with the CUDA library).  usage: _shard_worker.py <backend> <gname> <pe 0|1> <n> <out.npz>"""
import os
import sys

import numpy as np
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import common as cm
from bwa_mem_gpu_b200 import abi, lib as L, shard, synth


def main():
    backend, gname, pe, n, out = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
    dist.init_process_group("gloo")
    rank = dist.get_rank()
    lib = cm.backend_lib(backend)
    h = L.HostIndex(cm.genome(gname), lib)
    ctx = L.Context(rank if backend == "cuda" else 0, lib)
    ctx.upload_host_index(h)
    opt = abi.default_opt()
    pestat_fn = None
    if pe:
        opt.flag |= abi.MEM_F_PE
        reads, _ = synth.simulate_pairs(cm.genome(gname), n // 2, seed=29, damaged_frac=0.2, damage="periodic")
        seqs = np.ascontiguousarray(reads.reshape(-1))
        off = np.arange(0, n + 1, dtype=np.int64) * reads.shape[1]
        if rank == 0:                                   # the "host program" side: the reference's own mem_pestat
            idx = cm.ref_index(gname)
            pestat_fn = lambda regs, cnt: idx.pestat(opt, regs, cnt)
    else:
        seqs, off = cm.reads_for(gname, n, seed=47, indel=0.004)
    regs, cnt = shard.align_sharded(ctx, opt, seqs, off, dist=dist, pestat_fn=pestat_fn)
    if rank == 0:
        np.savez(out, regs=regs.view(np.uint8), cnt=cnt)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
