"""Config C4 (BASELINE.json: latent Gaussian model, train_lgm.py): S small scenes x V views at 512x512, forward+backward.
Three arms on the same GPU: the single-view drop-in API (one call per view, as train_lgm.py:111-120 does), the batched
entry point (one call per scene, row f1), all scenes in one call (gsr_forward_scenes) and - when oracle/_ref was built - the unmodified reference extension.
Prints one JSON line.  usage: python scripts/bench_c4.py [--scenes 8] [--iters 5]"""
import argparse, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "structured-gaussian-splatting_b200"), os.path.join(ROOT, "tests")]
import util  # noqa: E402
from gsr_b200 import synth, dp  # noqa: E402
import diff_gaussian_rasterization as dgr  # noqa: E402
from diff_gaussian_rasterization.batched import BatchedRasterizationSettings, GaussianRasterizerBatched  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scenes", type=int, default=8)
ap.add_argument("--iters", type=int, default=5)
args = ap.parse_args()
spec = synth.CONFIGS["C4"]
dev = torch.device("cuda", 0)
V = spec.V if hasattr(spec, "V") else 8
scenes = [dp.make_leaf_params(synth.spec_scene(spec, seed=s), dev) for s in range(args.scenes)]
cams = [synth.make_camera(k, V, spec.W, spec.H, spec.radius, spec.fovx_deg).to(dev) for k in range(V)]
bg = synth.spec_background(spec, dev)
go = synth.grad_image(spec.W, spec.H, device=dev)
go_b = go.unsqueeze(0).expand(V, -1, -1, -1).contiguous()
vm = torch.stack([c.viewmatrix for c in cams]); pm = torch.stack([c.projmatrix for c in cams]); cp = torch.stack([c.campos for c in cams])


def single(mod):
    for p in scenes:
        for c in cams:
            color, _, _ = dp.render_view(mod, p, c, bg, spec.D)
            color.backward(go)


def batched():
    rs = BatchedRasterizationSettings(spec.H, spec.W, cams[0].tanfovx, cams[0].tanfovy, bg, 1.0, vm, pm, spec.D, cp, False)
    for p in scenes:
        m2d = torch.zeros(V, spec.P, 3, device=dev, requires_grad=True)
        color, _ = GaussianRasterizerBatched(rs)(p["means3D"], m2d, p["opacities"], p["shs"], p["scales"], p["rotations"])
        color.backward(go_b)


# all scenes in ONE call (gsr_forward_scenes): parameters stacked [S,P,...], the same V cameras for every scene
stacked = {k: torch.stack([p[k].detach() for p in scenes]).requires_grad_(True) for k in scenes[0]}
S_ = args.scenes
vm_s = vm.unsqueeze(0).expand(S_, -1, -1, -1).contiguous(); pm_s = pm.unsqueeze(0).expand(S_, -1, -1, -1).contiguous()
cp_s = cp.unsqueeze(0).expand(S_, -1, -1).contiguous()
go_s = go_b.unsqueeze(0).expand(S_, -1, -1, -1, -1).contiguous()


def scenes_call():
    rs = BatchedRasterizationSettings(spec.H, spec.W, cams[0].tanfovx, cams[0].tanfovy, bg, 1.0, vm_s, pm_s, spec.D, cp_s, False)
    m2d = torch.zeros(S_, V, spec.P, 3, device=dev, requires_grad=True)
    color, _ = GaussianRasterizerBatched(rs)(stacked["means3D"], m2d, stacked["opacities"], stacked["shs"], stacked["scales"],
                                             stacked["rotations"])
    color.backward(go_s)
    for t in stacked.values():
        t.grad = None


def timeit(fn):
    for _ in range(2):
        fn()
    for p in scenes:
        for t in p.values():
            t.grad = None
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return args.scenes * V * args.iters / (e0.elapsed_time(e1) * 1e-3)


out = {"config": "C4: %d scenes x %d views, P=%d SH%d, %dx%d, fwd+bwd" % (args.scenes, V, spec.P, spec.D, spec.W, spec.H), "unit": "views/s"}
out["single_view_api"] = timeit(lambda: single(dgr))
out["batched_api"] = timeit(batched)
out["scenes_per_call_api"] = timeit(scenes_call)
# raw extension calls (no autograd wrapper) for an apples-to-apples comparison with the unmodified reference extension
raw_scenes = [{k: v.detach() for k, v in p.items()} for p in scenes]
cams_host = cams


def raw(C):
    for sc in raw_scenes:
        for c in cams_host:
            a, f = util.raw_forward(C, sc, c, bg, spec.D)
            util.raw_backward(C, a, f, go)


out["single_view_raw_calls"] = timeit(lambda: raw(dgr._C))
ref, _ = util.load_reference()
out["reference_extension_raw_calls"] = timeit(lambda: raw(ref)) if ref is not None else None
print(json.dumps(out))
