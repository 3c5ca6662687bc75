"""Multi-rank sharding (one process per GPU): contiguous blocks of reads per rank, index replicated,
results gathered on rank 0 in input order, batch-level mem_pestat computed once and broadcast.  The gathered result
must equal the reference's single-process result bit for bit."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import common as cm
from bwa_mem_gpu_b200 import abi, shard, synth

needs_ref = pytest.mark.skipif(not cm.have_ref(), reason="oracle/_ref not built")
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _run_world(world, backend, gname, pe, n, out):
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), GLOO_SOCKET_IFNAME="lo")
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "_shard_worker.py"), backend, gname, str(pe),
                                       str(n), out], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    logs = [p.communicate(timeout=900)[0].decode(errors="replace") for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(logs)


def test_block_bounds():
    assert shard.block_bounds(10, 3) == [(0, 3), (3, 6), (6, 10)]
    assert shard.block_bounds(10, 3, pair_aligned=True) == [(0, 2), (2, 6), (6, 10)]
    assert shard.block_bounds(0, 2) == [(0, 0), (0, 0)]
    for n in (1, 2, 7, 64, 1001):
        for w in (1, 2, 3, 8):
            b = shard.block_bounds(n - n % 2, w, pair_aligned=True)
            assert b[0][0] == 0 and b[-1][1] == n - n % 2 and all(x[1] == y[0] for x, y in zip(b, b[1:]))
            assert all(lo % 2 == 0 for lo, _ in b)


@needs_ref
@pytest.mark.parametrize("pe", [0, 1])
def test_two_ranks_gloo_emu(tmp_path, pe):
    gname, n = "small", 600
    out = str(tmp_path / "shard.npz")
    _run_world(2, "emu", gname, pe, n, out)
    g = np.load(out)
    idx = cm.ref_index(gname)
    opt = abi.default_opt()
    if pe:
        opt.flag |= abi.MEM_F_PE
        reads, _ = synth.simulate_pairs(cm.genome(gname), n // 2, seed=29, damaged_frac=0.2, damage="periodic")
        seqs = np.ascontiguousarray(reads.reshape(-1))
        off = np.arange(0, n + 1, dtype=np.int64) * reads.shape[1]
        regs, cnt = idx.align(opt, seqs, off, n_threads=8, cap_per_read=256)
        pes = idx.pestat(opt, regs, cnt)
        regs, cnt, n_sw = idx.mate_rescue(opt, pes, seqs, off, regs, cnt)
        assert n_sw > 0
    else:
        seqs, off = cm.reads_for(gname, n, seed=47, indel=0.004)
        regs, cnt = idx.align(opt, seqs, off, n_threads=8, cap_per_read=256)
    assert np.array_equal(g["cnt"], cnt)
    assert g["regs"].tobytes() == regs.tobytes()
