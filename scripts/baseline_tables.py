"""Measures the cells of BASELINE.md section 2 on the GPU box: the unmodified reference extension (oracle/_ref) and ours,
forward / backward ms and R per config (raw _C calls on one view, CUDA events), distCUDA2 at 3 M points, and the CPU
restatement (oracle/) on the box's host cores.  Prints one JSON object.
    python scripts/baseline_tables.py > gpurun_out/baseline_tables.json"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "structured-gaussian-splatting_b200"), os.path.join(ROOT, "tests")]
import util  # noqa: E402
import oracle  # noqa: E402
from gsr_b200 import synth  # noqa: E402
import diff_gaussian_rasterization as dgr  # noqa: E402
import simple_knn._C as knn  # noqa: E402

C = dgr._C
ref, refknn = util.load_reference()


def ev(fn, n=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n


out = {"cores": oracle.set_num_threads(len(os.sched_getaffinity(0))), "gpu": torch.cuda.get_device_name(0)}
for name, cam_k, V in (("C1", 0, 1), ("C3", 0, 64), ("C4", 3, 8), ("C5", 5, 64)):
    spec = synth.CONFIGS[name]
    sc = {k: v.cuda() for k, v in synth.spec_scene(spec).items()}
    cam = synth.make_camera(cam_k, V, spec.W, spec.H, spec.radius, spec.fovx_deg)
    bg = synth.spec_background(spec)
    go = synth.grad_image(spec.W, spec.H, seed=6).cuda()
    row = {"P": spec.P, "D": spec.D, "W": spec.W, "H": spec.H}
    for tag, mod in (("ours", C), ("ref", ref)):
        if mod is None:
            continue
        a, f = util.raw_forward(mod, sc, cam, bg, spec.D)
        row["R"] = int(f[0])
        row[tag + "_fwd_ms"] = round(ev(lambda: util.raw_forward(mod, sc, cam, bg, spec.D)), 4)
        row[tag + "_bwd_ms"] = round(ev(lambda: util.raw_backward(mod, a, f, go)), 4)
        del a, f
    out[name] = row
    print(name, row, file=sys.stderr, flush=True)
    if name == "C3":
        pts = sc["means3D"]
        out["dist2_3M"] = {"ours_ms": round(ev(lambda: knn.distCUDA2(pts), n=3, warm=1), 3)}
        if refknn is not None:
            out["dist2_3M"]["ref_ms"] = round(ev(lambda: refknn.distCUDA2(pts), n=2, warm=1), 3)
    del sc
    torch.cuda.empty_cache()

# CPU restatement: C1 forward and forward+backward (every tile); C3 from bench.py's cpu_baseline; kNN brute force at 1e5
spec = synth.CONFIGS["C1"]
sc = synth.spec_scene(spec); cam = synth.spec_camera(spec); bg = synth.spec_background(spec)
go = synth.grad_image(spec.W, spec.H, seed=6)
util.run_oracle(sc, cam, bg, spec.D, go)
t0 = time.perf_counter(); st, _ = util.run_oracle(sc, cam, bg, spec.D, None); t_f = time.perf_counter() - t0
t0 = time.perf_counter(); util.run_oracle(sc, cam, bg, spec.D, go); t_fb = time.perf_counter() - t0
out["cpu_C1"] = {"fwd_ms": round(t_f * 1e3, 2), "fwd_bwd_ms": round(t_fb * 1e3, 2)}
pts = np.random.default_rng(0).random((100000, 3), dtype=np.float32)
t0 = time.perf_counter(); oracle.dist2(pts, method="morton"); out["cpu_dist2_1e5_morton_ms"] = round((time.perf_counter() - t0) * 1e3, 1)
try:
    t0 = time.perf_counter(); oracle.dist2(pts[:20000], method="brute"); out["cpu_dist2_2e4_brute_ms"] = round((time.perf_counter() - t0) * 1e3, 1)
except Exception as ex:
    out["cpu_dist2_brute"] = str(ex)[:100]
print(json.dumps(out))
