"""Batched multi-view rasterization (extension of the reference API).

The reference renders one camera per call (`train_lgm.py:111-120` loops over views, `gaussian_renderer/__init__.py:18`).
`GaussianRasterizerBatched` renders V cameras over ONE Gaussian set in a single pass of the B200 pipeline: one
preprocess launch over (view, Gaussian), one depth sort / tile split / blending over the V stacked images, one host
wait instead of V.  Every view's image, radii and gradients equal what V single-view `GaussianRasterizer` calls give.

    settings = BatchedRasterizationSettings(image_height, image_width, tanfovx, tanfovy, bg, scale_modifier,
                                            viewmatrices[V,4,4], projmatrices[V,4,4], sh_degree, camposs[V,3], debug)
    colors[V,3,H,W], radii[V,P] = GaussianRasterizerBatched(settings)(means3D, means2D[V,P,3], opacities, shs, scales, rotations)

Parameter gradients come back summed over the views; `means2D.grad` stays per view (densification statistics).
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C
from . import _grad_target
import diff_gaussian_rasterization as _dgr


class BatchedRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrices: torch.Tensor     # [V,4,4], row-vector convention like GaussianRasterizationSettings.viewmatrix
    projmatrices: torch.Tensor     # [V,4,4]
    sh_degree: int
    camposs: torch.Tensor          # [V,3]
    debug: bool = False


class _RasterizeGaussiansBatched(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, opacities, scales, rotations, rs):
        R, color, radii, geom, binning, img = _C.rasterize_gaussians_batched(
            rs.bg, means3D, opacities, scales, rotations, rs.scale_modifier, rs.viewmatrices, rs.projmatrices, rs.tanfovx,
            rs.tanfovy, rs.image_height, rs.image_width, sh, rs.sh_degree, rs.camposs, rs.debug)
        ctx.rs = rs
        ctx.R = R
        # diff_gaussian_rasterization.set_grad_mode: write / add the summed gradients straight into the leaves' .grad
        ctx.grad_mode = _dgr._GRAD_MODE
        targets = tuple(_grad_target(t) for t in (means3D, opacities, sh, scales, rotations)) if ctx.grad_mode else ()
        ctx.grad_targets = targets if targets and all(t is not None for t in targets) else None
        ctx.save_for_backward(means3D, scales, rotations, radii, sh, geom, binning, img)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_color, _grad_radii):
        rs = ctx.rs
        means3D, scales, rotations, radii, sh, geom, binning, img = ctx.saved_tensors
        g2d, gop, g3d, gsh, gsc, grot = _C.rasterize_gaussians_batched_backward(
            rs.bg, means3D, radii, scales, rotations, rs.scale_modifier, rs.viewmatrices, rs.projmatrices, rs.tanfovx, rs.tanfovy,
            grad_color, sh, rs.sh_degree, rs.camposs, geom, ctx.R, binning, img, rs.debug,
            *(ctx.grad_targets if ctx.grad_targets else (torch.Tensor([]),) * 5), ctx.grad_mode == "accumulate")
        if ctx.grad_targets:
            return None, g2d, None, None, None, None, None
        return g3d, g2d, gsh, gop, gsc, grot, None


class GaussianRasterizerBatched(nn.Module):
    def __init__(self, raster_settings: BatchedRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, shs, scales, rotations):
        rs = self.raster_settings
        V = rs.viewmatrices.shape[0]
        if shs is None or scales is None or rotations is None:
            raise Exception('The batched rasterizer takes SHs, scales and rotations (no precomputed colours / covariances)')
        if means2D.shape[0] != V:
            raise Exception('means2D must be [V,P,3]: one screen-space gradient sink per view')
        return _RasterizeGaussiansBatched.apply(means3D, means2D, shs, opacities, scales, rotations, rs)
