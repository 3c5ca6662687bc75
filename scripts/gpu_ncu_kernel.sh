#!/bin/bash
# ncu --set full on kernels matching a regex, on a short bench run. Usage: scripts/gpu_ncu_kernel.sh <tag> <regex> [reads] [count]
TAG=$1; RE=$2; READS=${3:-1000000}; CNT=${4:-1}
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --reads $READS --steps 1 --warmup 1 --no-cpu-baseline"
timeout 600 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RE" -c $CNT -o $OUT/${TAG}_prof -f $CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "rc=$?"; tail -3 $OUT/${TAG}_ncu_full.log
