#!/bin/bash
# quick GPU sanity + tunable sweep. Usage: scripts/gpu_sweep.sh "<tune1>" "<tune2>" ...
OUT=gpurun_out; mkdir -p $OUT
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for T in "$@"; do
  echo "== tune: $T"
  timeout 400 python bench.py --reads 4000000 --steps 2 --warmup 1 --no-cpu-baseline --tune "$T" 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), d['stage_ms'], round(d['roofline']['frac'],3), d['extension']['gcups'])"
done
