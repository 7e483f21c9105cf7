"""Batched multi-view rasterization (extension of the reference API).

The reference renders one camera per call (`train_lgm.py:111-120` loops over views, `gaussian_renderer/__init__.py:18`).
`GaussianRasterizerBatched` renders V cameras over ONE Gaussian set in a single pass of the B200 pipeline: one
preprocess launch over (view, Gaussian), one depth sort / tile split / blending over the V stacked images, one host
wait instead of V.  Every view's image, radii and gradients equal what V single-view `GaussianRasterizer` calls give.

    settings = BatchedRasterizationSettings(image_height, image_width, tanfovx, tanfovy, bg, scale_modifier,
                                            viewmatrices[V,4,4], projmatrices[V,4,4], sh_degree, camposs[V,3], debug)
    colors[V,3,H,W], radii[V,P] = GaussianRasterizerBatched(settings)(means3D, means2D[V,P,3], opacities, shs, scales, rotations)

Parameter gradients come back summed over the views; `means2D.grad` stays per view (densification statistics).

Several scenes per call (config C4: many small scenes, a few views each): give every parameter a leading scene dimension
(means3D [S,P,3], opacities [S,P,1], shs [S,P,M,3], scales [S,P,3], rotations [S,P,4], means2D [S,Vs,P,3]) and the cameras as
viewmatrices / projmatrices [S,Vs,4,4], camposs [S,Vs,3]: colors come back as [S,Vs,3,H,W], radii as [S,Vs,P], each scene's
gradients summed over its own views.  Preprocess and the per-Gaussian backward run once per scene, everything in between
once over all scenes (gsr_forward_scenes / gsr_backward_scenes).
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
    def forward(ctx, means3D, means2D, sh, opacities, scales, rotations, rs, sh_rest):
        R, color, radii, geom, binning, img = _C.rasterize_gaussians_batched(
            rs.bg, means3D, opacities, scales, rotations, rs.scale_modifier, rs.viewmatrices, rs.projmatrices, rs.tanfovx,
            rs.tanfovy, rs.image_height, rs.image_width, sh, rs.sh_degree, rs.camposs, rs.debug, sh_rest)
        ctx.rs = rs
        ctx.R = R
        ctx.split = sh_rest.numel() != 0
        # diff_gaussian_rasterization.set_grad_mode: write / add the summed gradients straight into the leaves' .grad
        ctx.grad_mode = _dgr._GRAD_MODE
        leaves = (means3D, opacities, sh, scales, rotations) + ((sh_rest,) if ctx.split else ())
        targets = tuple(_grad_target(t) for t in leaves) if ctx.grad_mode else ()
        ctx.grad_targets = targets if targets and all(t is not None for t in targets) else None
        ctx.save_for_backward(means3D, scales, rotations, radii, sh, geom, binning, img, sh_rest)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_color, _grad_radii):
        rs = ctx.rs
        means3D, scales, rotations, radii, sh, geom, binning, img, sh_rest = ctx.saved_tensors
        e = torch.Tensor([])
        t = ctx.grad_targets
        into = (tuple(t[:5]) if t else (e,) * 5)
        into_rest = t[5] if (t and ctx.split) else e
        g2d, gop, g3d, gsh, gsc, grot, grest = _C.rasterize_gaussians_batched_backward(
            rs.bg, means3D, radii, scales, rotations, rs.scale_modifier, rs.viewmatrices, rs.projmatrices, rs.tanfovx, rs.tanfovy,
            grad_color, sh, rs.sh_degree, rs.camposs, geom, ctx.R, binning, img, rs.debug,
            *into, ctx.grad_mode == "accumulate", sh_rest, into_rest)
        if t:
            return None, g2d, None, None, None, None, None, None
        return g3d, g2d, gsh, gop, gsc, grot, None, (grest if ctx.split else None)


class GaussianRasterizerBatched(nn.Module):
    def __init__(self, raster_settings: BatchedRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, shs, scales, rotations, shs_rest=None):
        """shs_rest given: split SH storage - shs is the model's features_dc [P,1,3], shs_rest its features_rest [P,15,3]
        (read in place, see diff_gaussian_rasterization.split)."""
        rs = self.raster_settings
        V = rs.viewmatrices.shape[0]
        if shs is None or scales is None or rotations is None:
            raise Exception('The batched rasterizer takes SHs, scales and rotations (no precomputed colours / covariances)')
        if means3D.dim() == 3:
            # several scenes per call: every parameter carries a leading scene dimension S, the cameras are [S,Vs,...]
            if rs.viewmatrices.dim() != 4 or rs.viewmatrices.shape[0] != means3D.shape[0]:
                raise Exception('scenes: means3D [S,P,3] needs viewmatrices / projmatrices [S,Vs,4,4] and camposs [S,Vs,3]')
            if tuple(means2D.shape[:3]) != (means3D.shape[0], rs.viewmatrices.shape[1], means3D.shape[1]):
                raise Exception('scenes: means2D must be [S,Vs,P,3]')
        elif means2D.shape[0] != V:
            raise Exception('means2D must be [V,P,3]: one screen-space gradient sink per view')
        if shs_rest is not None and (shs.shape[-2] != 1 or shs_rest.dim() != shs.dim() or shs_rest.shape[-2] not in (0, 15)):
            raise Exception('split SH storage: shs [P,1,3] and shs_rest [P,15,3] (or [P,0,3]) expected')
        rest = torch.Tensor([]) if (shs_rest is None or shs_rest.numel() == 0) else shs_rest
        return _RasterizeGaussiansBatched.apply(means3D, means2D, shs, opacities, scales, rotations, rs, rest)
