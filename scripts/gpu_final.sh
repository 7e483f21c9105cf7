#!/usr/bin/env bash
# Round-end visit of one GPU box: the whole -m gpu suite, the default bench line, the reference arm, the launch list of the
# same command and one ncu --set full capture of the two blending kernels (each ncu pass only after its command exited 0
# without ncu).  usage: gpu_final.sh <tag>
set -u
mkdir -p gpurun_out
TAG=${1:-final}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/${TAG}_bench_reference_arm.json 2> gpurun_out/${TAG}_bench_reference_arm.err; echo "reference arm rc=$?"
python - <<PY
import json
for f in ("gpurun_out/${TAG}_bench.json", "gpurun_out/${TAG}_bench_reference_arm.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d.get("value"), d.get("unit"), "ms/view", d.get("ms_per_view"), "e2e", (d.get("e2e") or {}).get("value"))
    except Exception as e:
        print(f, "no result", e)
PY
bash scripts/gpu_launches.sh ${TAG}
timeout 300 ncu --set full --clock-control none --import-source on -k regex:blend_ -s 2 -c 2 -f -o gpurun_out/${TAG}_blend python scripts/one_view.py --iters 2 > gpurun_out/${TAG}_blend_ncu.log 2>&1; echo "ncu full rc=$?"
