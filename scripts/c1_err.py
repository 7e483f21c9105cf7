"""Gradient error of each blending variant against the committed reference fixture (config C1)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "structured-gaussian-splatting_b200"), os.path.join(ROOT, "tests")]
import util
from gsr_b200 import synth
import diff_gaussian_rasterization as dgr
C = dgr._C
g = np.load(os.path.join(ROOT, "tests", "golden", "ref_C1.npz"))
spec = synth.CONFIGS["C1"]; sc = synth.spec_scene(spec); cam = synth.spec_camera(spec); bg = synth.spec_background(spec)
go = synth.grad_image(cam.W, cam.H, seed=6)
names = ["means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations"]
for mode in sys.argv[1:] or ["1", "3", "4", "0"]:
    os.environ["GSR_BLEND"] = mode
    a0, f0 = util.raw_forward(C, sc, cam, bg, spec.D)
    g0 = util.raw_backward(C, a0, f0, go)
    out = []
    for n, x in zip(names, g0):
        if x.numel() == 0 or n in ("colors", "cov3D"):
            continue
        out.append("%s %.2e" % (n, util.rel_err(x.cpu().numpy(), g["grad_" + n])))
    print("GSR_BLEND=%s: %s" % (mode, "  ".join(out)), flush=True)
