#!/usr/bin/env bash
# ncu --set full of the deferred per-Gaussian backward kernels (one short bench step); plain run first as the recipe asks.
set -u
TAG=$1
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-extras --no-c5 --no-train-step"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"views_kernel|preprocess_bwd_kernel|preprocess_fwd|emit_instances" -s 12 -c 8 -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
