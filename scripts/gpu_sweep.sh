#!/bin/bash
# quick GPU sanity + tunable sweep. Usage: scripts/gpu_sweep.sh "<tune1>" "<tune2>" ...   ("-" = defaults)
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for T in "$@"; do
  echo "== tune: $T"
  TT=$T; [ "$T" = "-" ] && TT=""
  timeout 400 python bench.py --reads 4000000 --steps 2 --warmup 1 --no-cpu-baseline --tune "$TT" 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); c=d['counters']; print(round(d['value']), {k:round(v,1) for k,v in d['stage_ms'].items()}, round(d['roofline']['frac'],3), round(d['extension']['gcups'],1), [c.get('ext_bin0'),c.get('ext_bin1'),c.get('ext_bin2')])"
done
