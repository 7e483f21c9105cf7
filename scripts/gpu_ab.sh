#!/usr/bin/env bash
# A/B of kernel variants on one GPU box: parity tests, then short device-only bench runs under different env switches.
# usage: gpu_ab.sh <tag> "<ENV=..;ENV=..>" ...   (each further argument is one arm: a ';'-separated env list, '-' = defaults)
set -u
mkdir -p gpurun_out
TAG=$1; shift
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
tail -5 gpurun_out/${TAG}_pytest.log
i=0
for arm in "$@"; do
  i=$((i+1))
  envs=""
  if [ "$arm" != "-" ]; then envs=$(echo "$arm" | tr ';' ' '); fi
  echo "--- arm $i: $arm"
  env $envs python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${TAG}_arm${i}.json 2> gpurun_out/${TAG}_arm${i}.err
  echo "rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${TAG}_arm${i}.json").read().strip().splitlines()[-1])
    print("ms/view %.4f" % d["ms_per_view"], {k: v["ms"] for k, v in d["stages"].items()})
except Exception as e:
    print("no result", e)
PY
done
