"""One C3 view forward+backward a few times (GPU box): the command ncu captures are taken from.
    python scripts/one_view.py [--config C3] [--iters 3] [--opt name=value ...]"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "structured-gaussian-splatting_b200"), os.path.join(ROOT, "tests")]
import util  # noqa: E402
import gsr_b200  # noqa: E402
from gsr_b200 import synth  # noqa: E402
import diff_gaussian_rasterization as dgr  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="C3")
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--opt", action="append", default=[])
args = ap.parse_args()
C = dgr._C
opts = dict((kv.split("=")[0], int(kv.split("=")[1])) for kv in args.opt)
if any(k in ("blend_arm", "sort_cub", "blend_stats") for k in opts):
    gsr_b200.load_arms()
for k, v in opts.items():
    C.set_option(k, v)
spec = synth.CONFIGS[args.config]
sc = {k: v.cuda() for k, v in synth.spec_scene(spec).items()}
cam = synth.make_camera(0, 64, spec.W, spec.H, spec.radius, spec.fovx_deg)
bg = synth.spec_background(spec)
go = synth.grad_image(spec.W, spec.H, seed=6).cuda()
C.timing_reset(); C.timing_enable(True)
for _ in range(args.iters):
    a, f = util.raw_forward(C, sc, cam, bg, spec.D)
    g = util.raw_backward(C, a, f, go)
torch.cuda.synchronize()
print({name: round(ms / max(1, calls), 4) for name, ms, calls in C.timing_collect() if calls})
