#!/usr/bin/env bash
# ncu --set full capture of selected kernels of one short bench run (plain run first, as the recipe asks).
# usage: gpu_ncu_full.sh <tag> <kernel regex> <skip> <count>
set -u
TAG=$1; RE=$2; SKIP=$3; CNT=$4
CMD="python bench.py --steps 1 --warmup 3 --views-per-step 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$RE -s $SKIP -c $CNT -f -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
