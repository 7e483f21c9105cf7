"""C3 view, forward + backward from the model's two SH parameters: GaussianRasterizer over torch.cat((f_dc, f_rest)) (what
gaussian_renderer does through get_features) against GaussianRasterizerSplit reading them in place.  CUDA events, ms per view."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "structured-gaussian-splatting_b200"), ROOT]
import diff_gaussian_rasterization as dgr  # noqa: E402
from diff_gaussian_rasterization.split import GaussianRasterizerSplit  # noqa: E402
from gsr_b200 import synth  # noqa: E402


def main():
    spec = synth.CONFIGS[os.environ.get("CONFIG", "C3")]
    dev = torch.device("cuda")
    scene = synth.spec_scene(spec)
    cams = [synth.make_camera(k, 64, spec.W, spec.H, spec.radius, spec.fovx_deg).to(dev) for k in range(8)]
    bg = synth.spec_background(spec, dev)
    go = synth.grad_image(spec.W, spec.H, device=dev)
    leaf = lambda t: t.to(dev).clone().requires_grad_(True)
    p = {k: leaf(scene[k]) for k in ("means3D", "opacities", "scales", "rotations")}
    f_dc, f_rest = leaf(scene["shs"][:, :1].contiguous()), leaf(scene["shs"][:, 1:].contiguous())
    C = dgr._C

    def settings(cam):
        return dgr.GaussianRasterizationSettings(image_height=spec.H, image_width=spec.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                                                 bg=bg, scale_modifier=1.0, viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix,
                                                 sh_degree=spec.D, campos=cam.campos, prefiltered=False, debug=False)

    def zero():
        for t in list(p.values()) + [f_dc, f_rest]:
            t.grad = None

    def step_cat(k):
        zero()
        m2 = torch.zeros_like(p["means3D"], requires_grad=True)
        color, _ = dgr.GaussianRasterizer(settings(cams[k % 8]))(
            means3D=p["means3D"], means2D=m2, opacities=p["opacities"], shs=torch.cat((f_dc, f_rest), dim=1), scales=p["scales"],
            rotations=p["rotations"])
        color.backward(go)

    def step_split(k):
        zero()
        m2 = torch.zeros_like(p["means3D"], requires_grad=True)
        color, _ = GaussianRasterizerSplit(settings(cams[k % 8]))(p["means3D"], m2, p["opacities"], f_dc, f_rest, p["scales"],
                                                                   p["rotations"])
        color.backward(go)

    res = {"config": os.environ.get("CONFIG", "C3"), "P": int(p["means3D"].shape[0])}
    for name, fn in (("cat_ms_per_view", step_cat), ("split_ms_per_view", step_split)):
        for k in range(4):
            fn(k)
        C.timing_reset(); C.timing_enable(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        n = 16
        for k in range(n):
            fn(k)
        e1.record(); torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / n
        C.timing_enable(False)
        res[name.replace("_ms_per_view", "_stages")] = {s: round(ms / max(cnt, 1), 4) for s, ms, cnt in C.timing_collect() if cnt}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
