#!/usr/bin/env bash
# One GPU-box visit: parity tests, then the in-process step A/B (scripts/ab_step.py), then a short bench run (device + e2e legs).  usage: gpu_ab_step.sh <tag> [ab_step.py arguments]
set -u
mkdir -p gpurun_out
TAG=${1:-ab}; shift || true
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
tail -5 gpurun_out/${TAG}_pytest.log
python scripts/ab_step.py --out gpurun_out/${TAG}_ab_step.json "$@" > gpurun_out/${TAG}_ab_step.log 2>&1; echo "ab rc=$?"
tail -2 gpurun_out/${TAG}_ab_step.log
python bench.py --steps 8 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/${TAG}_bench_short.json 2> gpurun_out/${TAG}_bench_short.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${TAG}_bench_short.json").read().strip().splitlines()[-1])
    print("ms/view %.4f" % d["ms_per_view"], "e2e", d["e2e"]["value"] if d.get("e2e") else None, {k: v["ms"] for k, v in d["stages"].items()})
except Exception as e:
    print("no result", e)
PY
