"""A/B of the blending kernels on one C3 view (GPU box): per-stage device times of the product kernels and the comparison
arms, gradient agreement between them, and the statistics kernel's counters.  Not part of the product or the test-suite.
    python scripts/ab_blend.py [--config C3] [--iters 10]"""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "structured-gaussian-splatting_b200"), os.path.join(ROOT, "tests")]
import util  # noqa: E402
import gsr_b200  # noqa: E402
from gsr_b200 import synth  # noqa: E402
import diff_gaussian_rasterization as dgr  # noqa: E402

C = dgr._C


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C3")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--gaussians", type=int, default=0)
    args = ap.parse_args()
    spec = synth.CONFIGS[args.config]
    P = args.gaussians or spec.P
    sc = {k: v.cuda() for k, v in synth.spec_scene(spec, P=P).items()}
    cam = synth.make_camera(0, 64, spec.W, spec.H, spec.radius, spec.fovx_deg)
    bg = synth.spec_background(spec)
    go = synth.grad_image(spec.W, spec.H, seed=6).cuda()
    arms = gsr_b200.load_arms()
    out = {"config": args.config, "P": P}

    def run(n, **opts):
        prev = {k: C.set_option(k, v) for k, v in opts.items()}
        try:
            for _ in range(2):
                a, f = util.raw_forward(C, sc, cam, bg, spec.D)
                g = util.raw_backward(C, a, f, go)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                a, f = util.raw_forward(C, sc, cam, bg, spec.D)
                g = util.raw_backward(C, a, f, go)
            e1.record()
            torch.cuda.synchronize()
            total = e0.elapsed_time(e1) / n
            C.timing_reset(); C.timing_enable(True)
            for _ in range(n):
                a, f = util.raw_forward(C, sc, cam, bg, spec.D)
                g = util.raw_backward(C, a, f, go)
            torch.cuda.synchronize()
            C.timing_enable(False)
            rows = {name: round(ms / max(1, calls), 4) for name, ms, calls in C.timing_collect() if calls}
            rows["TOTAL_fwd_bwd"] = round(total, 4)
            return rows, f, g
        finally:
            for k, v in prev.items():
                C.set_option(k, v)

    variants = {
        "default": dict(),
        "bwd7_guarded": dict(bwd_variant=70),
        "fwd3_bwd6": dict(fwd_variant=3, bwd_variant=6),
        "default_again": dict(),
    }
    ref_g = None
    for name, kw in variants.items():
        try:
            rows, f, g = run(args.iters, **kw)
            entry = {"stages_ms": rows, "R": int(f[0])}
            if ref_g is None:
                ref_g, ref_f = g, f
            else:
                entry["image_equal"] = bool(torch.equal(f[1], ref_f[1]))
                entry["grad_rel_err_vs_first"] = max(util.rel_err(x.cpu().numpy(), y.cpu().numpy()) for x, y in zip(g, ref_g) if x.numel())
            out[name] = entry
        except Exception as ex:      # keep going: every arm is reported
            out[name] = {"error": str(ex)[:300]}
        print(name, json.dumps(out[name]), flush=True)

    # statistics kernel: what the culled forward evaluates / blends
    try:
        arms.gsr_debug_blend_stats(None, 1)
        C.set_option("blend_stats", 1)
        util.raw_forward(C, sc, cam, bg, spec.D)
        torch.cuda.synchronize()
        C.set_option("blend_stats", 0)
        st = (ctypes.c_ulonglong * 16)()
        arms.gsr_debug_blend_stats(st, 1)
        names = ["entries_staged", "warp_entry_pairs", "lanes_not_done", "lanes_power_le0", "lanes_alpha_ok", "lanes_blended",
                 "warp_entry_pairs_with_live_lane", "half_warp_pairs_live", "quarter_pairs_live", "batches"]
        out["stats"] = {n: int(st[i]) for i, n in enumerate(names)}
        print("stats", json.dumps(out["stats"]), flush=True)
    except Exception as ex:
        out["stats"] = {"error": str(ex)[:300]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
