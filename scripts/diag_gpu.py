"""First-contact diagnostics on the GPU box: stage-by-stage comparison of OUR path against the
reference extension (oracle/_ref) and rough timings, printed verbosely (pytest -x stops at the first
failure; this keeps going).  Not part of the product or the test-suite."""
import os
import sys
import time
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "structured-gaussian-splatting_b200"), os.path.join(ROOT, "tests")]
import util  # noqa: E402
from gsr_b200 import synth  # noqa: E402
import diff_gaussian_rasterization as dgr  # noqa: E402

C = dgr._C
ref, refknn = util.load_reference()
print("reference loaded:", ref is not None, flush=True)


def timeit(fn, n=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def compare(name, sc, cam, bg, D, timing=False):
    print("=== %s: P=%d D=%d %dx%d" % (name, sc["means3D"].shape[0], D, cam.W, cam.H), flush=True)
    P = sc["means3D"].shape[0]
    scd = {k: v.cuda() for k, v in sc.items()}
    try:
        a0, f0 = util.raw_forward(C, scd, cam, bg, D)
        torch.cuda.synchronize()
        print("ours: R=%d visible=%d" % (f0[0], int((f0[2] > 0).sum())), flush=True)
        e0 = util.export_ours(C, f0, P, cam.W, cam.H)
        go = synth.grad_image(cam.W, cam.H, seed=6)
        g0 = util.raw_backward(C, a0, f0, go)
        torch.cuda.synchronize()
        if ref is not None:
            a1, f1 = util.raw_forward(ref, scd, cam, bg, D)
            torch.cuda.synchronize()
            print("ref : R=%d visible=%d" % (f1[0], int((f1[2] > 0).sum())))
            print("radii mismatches:", int((f0[2] != f1[2]).sum()))
            if f0[0] == f1[0]:
                e1 = util.parse_reference_blobs(f1[4], f1[5], f1[0], cam.W, cam.H)
                for k in ("vals", "keys", "ranges", "n_contrib"):
                    print("  %-10s mismatches: %d / %d" % (k, int((e0[k] != e1[k]).sum()), e0[k].size))
                print("  final_T max abs diff:", float(np.abs(e0["final_T"] - e1["final_T"]).max()))
            d = (f0[1] - f1[1]).abs()
            print("image max abs diff: %.3e  (#pix > 1e-6: %d, exact equal: %s)" % (float(d.max()), int((d > 1e-6).sum()), bool(torch.equal(f0[1], f1[1]))))
            g1 = util.raw_backward(ref, a1, f1, go)
            torch.cuda.synchronize()
            for n, x, y in zip(["means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations"], g0, g1):
                if x.numel() == 0: continue
                print("  grad %-10s rel err %.3e (max ref %.3e)" % (n, util.rel_err(x.cpu().numpy(), y.cpu().numpy()), float(y.abs().max())))
        if timing:
            go_d = go.cuda()
            t_f = timeit(lambda: util.raw_forward(C, scd, cam, bg, D))
            fwd = util.raw_forward(C, scd, cam, bg, D)
            t_b = timeit(lambda: util.raw_backward(C, fwd[0], fwd[1], go_d))
            print("ours fwd %.3f ms  bwd %.3f ms" % (t_f, t_b))
            os.environ["GSR_SORT"] = "1"
            t_fc = timeit(lambda: util.raw_forward(C, scd, cam, bg, D))
            os.environ["GSR_SORT"] = "0"
            print("ours fwd with cub arm %.3f ms" % t_fc)
            if ref is not None:
                t_f = timeit(lambda: util.raw_forward(ref, scd, cam, bg, D))
                fwd = util.raw_forward(ref, scd, cam, bg, D)
                t_b = timeit(lambda: util.raw_backward(ref, fwd[0], fwd[1], go_d))
                print("ref  fwd %.3f ms  bwd %.3f ms" % (t_f, t_b))
    except Exception:
        traceback.print_exc()
    sys.stdout.flush()


if __name__ == "__main__":
    which = sys.argv[1:] or ["small", "C1", "C4", "C2"]
    if "small" in which:
        sc, cam, bg = util.small_scene(5000, 3, 320, 200, 29)
        compare("small3", sc, cam, bg, 3)
    for name in ("C1", "C4", "C2"):
        if name in which:
            spec = synth.CONFIGS[name]
            compare(name, synth.spec_scene(spec), synth.spec_camera(spec), synth.spec_background(spec), spec.D, timing=True)
