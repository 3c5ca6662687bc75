#!/bin/bash
# Run on the GPU box (under gpurun): full-size bench (default flags), then the ncu launch list of a shorter run of the
# same command.  One ncu pass per call; the --set full capture is scripts/gpu_ncu_kernel.sh in its own call.
# Outputs under gpurun_out/<tag>_*.  Usage: scripts/gpu_profile.sh <tag> [reads_for_ncu]
set -u
TAG=${1:-r1}
NCU_READS=${2:-2000000}
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench rc=$?"; tail -c 3500 $OUT/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err
echo "ref rc=$?"; tail -c 600 $OUT/${TAG}_bench_ref.json
CMD="python bench.py --reads $NCU_READS --steps 1 --warmup 1 --no-cpu-baseline"
timeout 600 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_list.log 2>&1
echo "ncu list rc=$?"
