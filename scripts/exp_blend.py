"""Experiment harness (GPU box): blending-kernel variants selected with GSR_BLEND / GSR_BLEND_BWD on the C2 scene.
Checks every variant against variant 2 (image / n_contrib / final_T bit-equality, gradient relative error) and times
forward and backward with CUDA events.  Also prints the work statistics of blend_stats_kernel.  Not part of the product."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "structured-gaussian-splatting_b200"), os.path.join(ROOT, "tests")]
import util  # noqa: E402
from gsr_b200 import synth, LIB_PATH  # noqa: E402
import diff_gaussian_rasterization as dgr  # noqa: E402

C = dgr._C
lib = ctypes.CDLL(LIB_PATH)


def stage_ms(fn, n=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    C.timing_reset(); C.timing_enable(True)
    for _ in range(n): fn()
    torch.cuda.synchronize()
    C.timing_enable(False)
    return {name: ms / max(1, calls) for name, ms, calls in C.timing_collect()}


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
    fvars = [int(x) for x in (sys.argv[2].split(",") if len(sys.argv) > 2 else ["2"])]
    bvars = [int(x) for x in (sys.argv[3].split(",") if len(sys.argv) > 3 else ["2"])]
    spec = synth.CONFIGS[cfg]
    sc = {k: v.cuda() for k, v in synth.spec_scene(spec).items()}
    cam = synth.spec_camera(spec); bg = synth.spec_background(spec)
    go = synth.grad_image(spec.W, spec.H, seed=6).cuda()
    P = sc["means3D"].shape[0]

    os.environ["GSR_BLEND_STATS"] = "1"
    lib.gsr_debug_blend_stats(None, 1)
    util.raw_forward(C, sc, cam, bg, spec.D); torch.cuda.synchronize()
    os.environ["GSR_BLEND_STATS"] = "0"
    st = (ctypes.c_ulonglong * 16)()
    lib.gsr_debug_blend_stats(st, 1)
    names = ["entries_staged", "warp_entries", "lanes_active", "lanes_power_ok", "lanes_alpha_ok", "lanes_blended",
             "warp_entries_exact", "half_entries_exact", "quarter_entries_exact", "batches",
             "iters_two_4x4", "iters_two_8x2", "iters_four_4x2", "block44_entries_exact",
             "sum_batch_max_warp", "sum_tile_max_warp"]
    print("stats:", {n: int(st[i]) for i, n in enumerate(names)}, flush=True)

    base = None
    for v in fvars:
        os.environ["GSR_BLEND"] = str(v)
        a, f = util.raw_forward(C, sc, cam, bg, spec.D)
        ex = util.export_ours(C, f, P, cam.W, cam.H)
        img = f[1].clone()
        t = stage_ms(lambda: util.raw_forward(C, sc, cam, bg, spec.D))
        msg = "fwd variant %d: blend_forward %.4f ms (pre %.3f depth %.3f split %.3f)" % (
            v, t["blend_forward"], t["preprocess"], t["depth_order"], t["tile_split"])
        if base is None:
            base = (img, ex)
        else:
            msg += "  image equal=%s maxdiff=%.2e n_contrib equal=%s final_T equal=%s" % (
                bool(torch.equal(img, base[0])), float((img - base[0]).abs().max()),
                bool(np.array_equal(ex["n_contrib"], base[1]["n_contrib"])), bool(np.array_equal(ex["final_T"], base[1]["final_T"])))
        print(msg, flush=True)
    os.environ["GSR_BLEND"] = "0"
    a, f = util.raw_forward(C, sc, cam, bg, spec.D)
    gbase = None
    for v in bvars:
        os.environ["GSR_BLEND"] = str(v)
        g = [x.clone() for x in util.raw_backward(C, a, f, go)]
        t = stage_ms(lambda: util.raw_backward(C, a, f, go))
        msg = "bwd variant %d: blend_backward %.4f ms  preprocess_backward %.4f ms" % (v, t["blend_backward"], t["preprocess_backward"])
        if gbase is None:
            gbase = g
        else:
            errs = [util.rel_err(x.cpu().numpy(), y.cpu().numpy()) for x, y in zip(g, gbase) if x.numel()]
            msg += "  grad rel err vs first: max %.2e" % max(errs)
        print(msg, flush=True)


if __name__ == "__main__":
    main()
