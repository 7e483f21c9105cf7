"""Full training iteration at C3 (3 M SH-3 Gaussians, 1080p): gsr_b200.train_step.fused_step against reference_step (the
reference's torch formulation of everything around the rasterizer, over the same drop-in rasterizer).  ms per iteration."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "structured-gaussian-splatting_b200"), ROOT]
from gsr_b200 import synth, train_step  # noqa: E402


def main():
    spec = synth.CONFIGS[os.environ.get("CONFIG", "C3")]
    dev = torch.device("cuda")
    scene = synth.spec_scene(spec)
    cams = [synth.make_camera(k, 64, spec.W, spec.H, spec.radius, spec.fovx_deg).to(dev) for k in range(8)]
    bg = synth.spec_background(spec, dev)
    g = torch.Generator().manual_seed(3)
    gts = [torch.rand(3, spec.H, spec.W, generator=g).to(dev) for _ in range(2)]
    res = {"config": os.environ.get("CONFIG", "C3"), "P": int(scene["means3D"].shape[0])}
    for name, fused in (("reference_formulation_ms", False), ("fused_ms", True)):
        leaf = lambda t: t.to(dev).clone().contiguous().requires_grad_(True)
        model = {"xyz": leaf(scene["means3D"]), "features_dc": leaf(scene["shs"][:, :1]), "features_rest": leaf(scene["shs"][:, 1:]),
                 "scaling": leaf(torch.log(scene["scales"])), "rotation": leaf(scene["rotations"] * 1.7),
                 "opacity": leaf(torch.logit(scene["opacities"].clamp(1e-4, 1 - 1e-4)))}
        stats = train_step.new_stats(model["xyz"].shape[0], dev)
        opt = train_step.make_optimizer(model, fused)
        fn = train_step.fused_step if fused else train_step.reference_step
        for k in range(3):
            fn(model, stats, opt, cams[k % 8], bg, gts[k % 2], spec.D)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        n = 12
        for k in range(n):
            loss = fn(model, stats, opt, cams[k % 8], bg, gts[k % 2], spec.D)
        e1.record(); torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / n
        res[name.replace("_ms", "_loss")] = float(loss)
        del model, stats, opt
        torch.cuda.empty_cache()
    res["speedup"] = res["reference_formulation_ms"] / res["fused_ms"]
    print(json.dumps(res))


if __name__ == "__main__":
    main()
