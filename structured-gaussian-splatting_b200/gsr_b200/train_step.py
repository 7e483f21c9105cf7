"""One training iteration around the rasterizer, assembled from the fused pieces.

`fused_step` is the body of the reference's training loop (train.py:95-140) for one camera:
    render (gaussian_renderer/__init__.py:18-100: activations, get_features, rasterizer)
    loss = (1 - l) L1 + l (1 - SSIM); loss.backward()                      (train.py:102-106)
    max_radii2D / add_densification_stats                                   (train.py:129-132)
    optimizer.step(); optimizer.zero_grad(set_to_none=True)                 (train.py:138-140)
with  activate() -> GaussianRasterizerSplit -> l1_ssim_loss -> densify_stats -> FusedAdam.
`reference_step` is the same iteration written the way the reference writes it (torch activations, torch.cat, torch loss,
indexed statistics, torch.optim.Adam) over the drop-in GaussianRasterizer: the comparison arm for tests and
scripts/bench_train_step.py.  Both take a `model` dict of leaf parameters
    xyz [P,3], features_dc [P,1,3], features_rest [P,15,3], scaling [P,3], rotation [P,4], opacity [P,1]
and a `stats` dict max_radii2D [P], xyz_gradient_accum [P,1], denom [P,1].
"""
from __future__ import annotations

import torch

PARAM_LR = (("xyz", 0.00016), ("features_dc", 0.0025), ("features_rest", 0.0025 / 20.0), ("opacity", 0.05), ("scaling", 0.005),
            ("rotation", 0.001))                                            # arguments/__init__.py:79-86


def make_optimizer(model, fused: bool, spatial_lr_scale: float = 1.0):
    """the six groups of GaussianModel.training_setup (scene/gaussian_model.py:164-175)"""
    groups = [{"params": [model[n]], "lr": lr * (spatial_lr_scale if n == "xyz" else 1.0), "name": n} for n, lr in PARAM_LR]
    if fused:
        from .optim import FusedAdam
        return FusedAdam(groups, lr=0.0, eps=1e-15)
    return torch.optim.Adam(groups, lr=0.0, eps=1e-15)


def new_stats(P, device):
    return {"max_radii2D": torch.zeros(P, device=device), "xyz_gradient_accum": torch.zeros(P, 1, device=device),
            "denom": torch.zeros(P, 1, device=device)}


def _settings(dgr, cam, bg, sh_degree):
    return dgr.GaussianRasterizationSettings(image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=bg,
                                             scale_modifier=1.0, viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix,
                                             sh_degree=sh_degree, campos=cam.campos, prefiltered=False, debug=False)


def fused_step(model, stats, optimizer, cam, bg, gt_image, sh_degree: int, lambda_dssim: float = 0.2):
    import diff_gaussian_rasterization as dgr
    from diff_gaussian_rasterization.split import GaussianRasterizerSplit
    from .activations import activate
    from .losses import l1_ssim_loss
    from .optim import densify_stats
    scaling, rotation, opacity = activate(model["scaling"], model["rotation"], model["opacity"])
    means2D = torch.zeros_like(model["xyz"], requires_grad=True)
    image, radii = GaussianRasterizerSplit(_settings(dgr, cam, bg, sh_degree))(
        model["xyz"], means2D, opacity, model["features_dc"], model["features_rest"], scaling, rotation)
    loss = l1_ssim_loss(image, gt_image, lambda_dssim)
    loss.backward()
    with torch.no_grad():
        densify_stats(radii, means2D.grad, stats["max_radii2D"], stats["xyz_gradient_accum"], stats["denom"])
        optimizer.step()
        optimizer.zero_grad(set_to_none=True)
    return loss.detach()


def _torch_ssim(x, y):
    from math import exp
    g = torch.tensor([exp(-(i - 5) ** 2 / float(2 * 1.5 ** 2)) for i in range(11)], device=x.device)
    g = g / g.sum()
    C = x.shape[0]
    win = torch.outer(g, g)[None, None].expand(C, 1, 11, 11).contiguous()
    conv = lambda t: torch.nn.functional.conv2d(t[None], win, padding=5, groups=C)[0]
    mu1, mu2 = conv(x), conv(y)
    s1, s2, s12 = conv(x * x) - mu1 * mu1, conv(y * y) - mu2 * mu2, conv(x * y) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 0.01 ** 2) * (2 * s12 + 0.03 ** 2)) / ((mu1 * mu1 + mu2 * mu2 + 0.01 ** 2) * (s1 + s2 + 0.03 ** 2))
    return m.mean()


def reference_step(model, stats, optimizer, cam, bg, gt_image, sh_degree: int, lambda_dssim: float = 0.2):
    import diff_gaussian_rasterization as dgr
    scaling, rotation, opacity = torch.exp(model["scaling"]), torch.nn.functional.normalize(model["rotation"]), \
        torch.sigmoid(model["opacity"])
    shs = torch.cat((model["features_dc"], model["features_rest"]), dim=1)
    means2D = torch.zeros_like(model["xyz"], requires_grad=True)
    image, radii = dgr.GaussianRasterizer(_settings(dgr, cam, bg, sh_degree))(
        means3D=model["xyz"], means2D=means2D, opacities=opacity, shs=shs, scales=scaling, rotations=rotation)
    l1 = torch.abs(image - gt_image).mean()
    loss = (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - _torch_ssim(image, gt_image))
    loss.backward()
    with torch.no_grad():
        vis = radii > 0
        stats["max_radii2D"][vis] = torch.max(stats["max_radii2D"][vis], radii[vis].float())
        stats["xyz_gradient_accum"][vis] += torch.norm(means2D.grad[vis, :2], dim=-1, keepdim=True)
        stats["denom"][vis] += 1
        optimizer.step()
        optimizer.zero_grad(set_to_none=True)
    return loss.detach()
