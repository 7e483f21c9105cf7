#!/usr/bin/env bash
# One GPU-box visit: parity tests, bench, diagnostics vs the reference .so, ncu launch list (each ncu pass only after
# its command exited 0 without ncu).  Outputs land in gpurun_out/.
set -u
mkdir -p gpurun_out
TAG=${1:-x}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
python scripts/diag_gpu.py > gpurun_out/${TAG}_diag.log 2>&1; echo "diag rc=$?"
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; rc=$?; echo "bench rc=$rc"
cat gpurun_out/${TAG}_bench.json
if [ $rc -eq 0 ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv \
      python bench.py --steps 1 --warmup 3 --views-per-step 1 --no-cpu-baseline --no-e2e > gpurun_out/${TAG}_ncu.log 2>&1
  echo "ncu rc=$?"
fi
