#!/bin/bash
# Run on the GPU box (under gpurun): full-size bench, then the ncu launch list and one full capture of the
# two dominant kernels.  Outputs under gpurun_out/<tag>_*.  Usage: scripts/gpu_profile.sh <tag> [reads_for_ncu]
set -u
TAG=${1:-r1}
NCU_READS=${2:-2000000}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > $OUT/${TAG}_clocks.csv &
SMI=$!
timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench rc=$?"; tail -c 3000 $OUT/${TAG}_bench.json
kill $SMI
CMD="python bench.py --reads $NCU_READS --steps 1 --warmup 1 --no-cpu-baseline"
timeout 600 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'smem_collect_kernel|^extend_kernel|chain_kernel' -c 3 -o $OUT/${TAG}_prof -f $CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT
