#!/usr/bin/env bash
# ncu launch list (per-launch device time) of one short bench run; plain run first as the recipe asks.
set -u
TAG=$1
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-extras --no-c5 --no-train-step"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
