#!/usr/bin/env python
"""Measure the roofline denominators of the two hot stages on this GPU (writes JSON to stdout).
gather: dependent random 32-byte sector gathers/s (FM-index seeding pattern) for a 100 MB and a 3.1 GB table
dpx   : DPX instructions/s (ksw_extend2 cell arithmetic)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bwa_mem_gpu_b200 import lib as L
ctx = L.Context(0)
res = {"gather": [], "dpx": []}
for tb in (100 << 20, 3100 << 20):
    for bps in (4, 6, 8, 12, 16):
        for mlp in (1, 2, 4):
            ms, units = ctx.microbench(0, tb, bps, 1500, mlp)
            res["gather"].append({"table_mb": tb >> 20, "blocks_per_sm": bps, "mlp": mlp, "ms": ms,
                                  "gsectors_per_s": units / ms / 1e6, "gbytes_per_s": units * 32 / ms / 1e6})
for bps in (2, 4, 8):
    ms, units = ctx.microbench(1, 0, bps, 20000, 1)
    res["dpx"].append({"blocks_per_sm": bps, "ms": ms, "tera_dpx_per_s": units / ms / 1e9})
best = {}
for g in res["gather"]:
    k = str(g["table_mb"])
    if k not in best or g["gsectors_per_s"] > best[k]["gsectors_per_s"]:
        best[k] = g
res["best_gather"] = best
res["best_dpx"] = max(res["dpx"], key=lambda d: d["tera_dpx_per_s"])
print(json.dumps(res, indent=1))
