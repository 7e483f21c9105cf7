#!/usr/bin/env bash
# ncu launch list (per-launch device time) of one short bench run; plain run first as the recipe asks.
set -u
TAG=$1
CMD="python bench.py --steps 1 --warmup 3 --views-per-step 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
