"""Small cases for compute-sanitizer (GPU box): C1 forward+backward through the drop-in API, a batched 3-view call, a
two-range staged backward, distCUDA2 on 20 k points.  One tool per gpurun call:
    compute-sanitizer --tool racecheck|initcheck|memcheck|synccheck python scripts/sanitize_case.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "structured-gaussian-splatting_b200"), os.path.join(ROOT, "tests")]
import util  # noqa: E402
from gsr_b200 import synth, dp  # noqa: E402
import diff_gaussian_rasterization as dgr  # noqa: E402
from diff_gaussian_rasterization.batched import BatchedRasterizationSettings, GaussianRasterizerBatched  # noqa: E402
import simple_knn._C as knn  # noqa: E402

spec = synth.CONFIGS["C1"]
sc = synth.spec_scene(spec)
cam = synth.spec_camera(spec)
bg = synth.spec_background(spec)
go = synth.grad_image(cam.W, cam.H, seed=6)
r = util.run_ours(sc, cam, bg, spec.D, go)
print("C1 fwd+bwd ok", float(r.color.sum()), flush=True)

P, D, W, H, V = 4000, 3, 160, 120, 3
sc2, _, bg2 = util.small_scene(P, D, W, H, 43)
cams = [synth.make_camera(k, 5, W, H, 3.0, 50.0).to("cuda") for k in range(V)]
p2 = dp.make_leaf_params(sc2, "cuda")
rs = BatchedRasterizationSettings(H, W, cams[0].tanfovx, cams[0].tanfovy, bg2.cuda(), 1.0, torch.stack([c.viewmatrix for c in cams]),
                                  torch.stack([c.projmatrix for c in cams]), D, torch.stack([c.campos for c in cams]), False)
m2 = torch.zeros(V, P, 3, device="cuda", requires_grad=True)
color, radii = GaussianRasterizerBatched(rs)(p2["means3D"], m2, p2["opacities"], p2["shs"], p2["scales"], p2["rotations"])
color.sum().backward()
print("batched ok", float(color.sum()), flush=True)

p3 = dp.make_leaf_params(sc2, "cuda")
flat = dp.FlatGrads(p3)
prev = dgr.set_grad_mode("overwrite"); hook = dgr.set_chunk_hook((2, lambda a, b: None))
c3, _, _ = dp.render_view(dgr, p3, cams[0], bg2.cuda(), D)
c3.sum().backward()
dgr.set_grad_mode(prev); dgr.set_chunk_hook(hook)
print("staged ok", float(flat.flat.abs().sum()), flush=True)

pts = torch.rand(20000, 3, generator=torch.Generator().manual_seed(5)).cuda()
d = knn.distCUDA2(pts)
torch.cuda.synchronize()
print("dist2 ok", float(d.mean()), flush=True)
