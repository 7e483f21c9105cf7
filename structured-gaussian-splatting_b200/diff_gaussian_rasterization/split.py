"""Rasterization straight from the model's two SH parameters (extension of the reference API).

The reference model keeps `_features_dc [P,1,3]` and `_features_rest [P,15,3]` as separate parameters (different learning
rates, scene/gaussian_model.py:151-152,168-169) and concatenates them for every render (`get_features`, :119-122:
a 1.15 GB read+write at 3 M Gaussians) - and autograd splits the 576 MB SH gradient back into the two on the way down.
`GaussianRasterizerSplit` takes the two tensors as they are: the preprocess kernel reads the DC triple and the 45-float
row in place, the SH backward stages both through shared memory and writes `dL/dfeatures_dc` and `dL/dfeatures_rest`
directly.  Results are bit-identical to `GaussianRasterizer(shs=torch.cat((features_dc, features_rest), dim=1))`.

    color, radii = GaussianRasterizerSplit(raster_settings)(means3D, means2D, opacities, features_dc, features_rest,
                                                            scales, rotations)

`features_rest` must hold 15 coefficients (max_sh_degree 3) or none ([P,0,3]: plain DC colour); other sizes raise, use the
reference API with torch.cat for them.  diff_gaussian_rasterization.set_grad_mode applies (targets: the five parameter
gradients plus features_rest's).
"""
import torch
import torch.nn as nn

from . import _C
from . import _grad_target
import diff_gaussian_rasterization as _dgr


class _RasterizeGaussiansSplit(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, f_dc, f_rest, opacities, scales, rotations, rs):
        e = torch.Tensor([])
        R, color, radii, geom, binning, img = _C.rasterize_gaussians_split(
            rs.bg, means3D, e, opacities, scales, rotations, rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
            rs.tanfovy, rs.image_height, rs.image_width, f_dc, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug, f_rest)
        ctx.rs = rs
        ctx.R = R
        ctx.grad_mode = _dgr._GRAD_MODE
        ctx.grad_targets = (tuple(_grad_target(t) for t in (means3D, opacities, f_dc, scales, rotations, f_rest))
                            if ctx.grad_mode else None)
        ctx.save_for_backward(means3D, scales, rotations, radii, f_dc, f_rest, geom, binning, img)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_color, _grad_radii):
        rs = ctx.rs
        means3D, scales, rotations, radii, f_dc, f_rest, geom, binning, img = ctx.saved_tensors
        e = torch.Tensor([])
        t = ctx.grad_targets
        if t is not None and f_rest.numel() and (t[2] is None) != (t[5] is None):
            t = t[:2] + (None,) + t[3:5] + (None,)            # SH targets go together or not at all
        into = tuple(e if x is None else x for x in t) if t is not None else (e,) * 6
        g2d, _gcol, gop, g3d, _gcov, gdc, gsc, grot, grest = _C.rasterize_gaussians_backward_split(
            rs.bg, means3D, radii, e, scales, rotations, rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
            rs.tanfovy, grad_color, f_dc, rs.sh_degree, rs.campos, geom, ctx.R, binning, img, rs.debug,
            into[0], into[1], into[2], into[3], into[4], ctx.grad_mode == "accumulate", f_rest, into[5])
        if t is not None:
            g3d = None if t[0] is not None else g3d
            gop = None if t[1] is not None else gop
            gdc = None if t[2] is not None else gdc
            gsc = None if t[3] is not None else gsc
            grot = None if t[4] is not None else grot
            grest = None if t[5] is not None else grest
        if f_rest.numel() == 0:
            grest = None if (t is not None and t[5] is not None) else torch.zeros_like(f_rest)
        return g3d, g2d, gdc, grest, gop, gsc, grot, None


class GaussianRasterizerSplit(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, features_dc, features_rest, scales, rotations):
        if features_dc.dim() != 3 or features_dc.shape[1] != 1 or features_rest.dim() != 3:
            raise Exception("features_dc [P,1,3] and features_rest [P,K,3] expected")
        if features_rest.shape[1] not in (0, 15):
            raise Exception("GaussianRasterizerSplit supports 15 extra SH coefficients (or none); "
                            "use GaussianRasterizer(shs=torch.cat(...)) for other sizes")
        return _RasterizeGaussiansSplit.apply(means3D, means2D, features_dc, features_rest, opacities, scales, rotations,
                                              self.raster_settings)
