#!/bin/bash
# A/B of context tunables on the GPU box: scripts/gpu_ab.sh <tag> "<tune1>" "<tune2>" ...   (each a --tune string; "-" = defaults)
# One short bench per setting (4 M reads, 2 timed steps, no CPU baseline); stage times go to gpurun_out/<tag>_ab.txt
TAG=$1; shift
OUT=gpurun_out/${TAG}_ab.txt
mkdir -p gpurun_out; : > $OUT
for T in "$@"; do
  ARG=""; [ "$T" != "-" ] && ARG="--tune $T"
  timeout 300 python bench.py --reads 4000000 --steps 2 --warmup 2 --no-cpu-baseline $ARG $AB_EXTRA > gpurun_out/_ab.json 2> gpurun_out/_ab.err || { echo "$T FAILED" >> $OUT; tail -3 gpurun_out/_ab.err >> $OUT; continue; }
  python - "$T" >> $OUT <<'P'
import json,sys
d=json.load(open('gpurun_out/_ab.json'))
s=d['stage_ms']
print(f"{sys.argv[1]:40s} reads/s {d['value']/1e6:7.2f}M e2e {d['e2e']['value']/1e6:7.2f}M total {s['total']:7.2f} smem {s['smem']:7.2f} ext {s['extend']:7.2f} chain {s['chain']:5.2f} sa {s['seed_sa']:5.2f} gcups {d['extension']['gcups']:6.1f} truth {d['truth_check']['frac']:.4f}")
P
done
cat $OUT
