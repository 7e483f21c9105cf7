"""Shared helpers of the parity tests: run OUR extension, the CPU oracle, and (when present) the
unmodified reference CUDA extension built into oracle/_ref, on the same seeded inputs."""
from __future__ import annotations

import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")

from gsr_b200 import synth  # noqa: E402


def small_scene(P=2000, D=1, W=160, H=120, seed=0, ln_scale=math.log(0.03), half=1.0, radius=3.0, fov=50.0):
    sc = synth.make_scene(P, D, half, ln_scale, seed)
    cam = synth.make_camera(0, 1, W, H, radius, fov)
    bg = torch.rand(3, generator=torch.Generator().manual_seed(1))
    return sc, cam, bg


def run_oracle(sc, cam, bg, D, grad_out=None, use_sh=True, precomp_cov=None, precomp_col=None, scale_modifier=1.0):
    import oracle
    kw = dict(viewmatrix=cam.viewmatrix.numpy(), projmatrix=cam.projmatrix.numpy(), campos=cam.campos.numpy(),
              bg=bg.numpy(), W=cam.W, H=cam.H, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, sh_degree=D,
              scale_modifier=scale_modifier)
    if precomp_col is not None:
        kw["colors_precomp"] = precomp_col
    else:
        kw["shs"] = sc["shs"].numpy()
    if precomp_cov is not None:
        kw["cov3D_precomp"] = precomp_cov
    else:
        kw["scales"] = sc["scales"].numpy(); kw["rotations"] = sc["rotations"].numpy()
    st = oracle.forward(sc["means3D"].numpy(), sc["opacities"].numpy(), **kw)
    grads = oracle.backward(st, grad_out.numpy()) if grad_out is not None else None
    return st, grads


def settings(mod, cam, bg, D, dev, scale_modifier=1.0, debug=False):
    return mod.GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=bg.to(dev),
        scale_modifier=scale_modifier, viewmatrix=cam.viewmatrix.to(dev), projmatrix=cam.projmatrix.to(dev), sh_degree=D,
        campos=cam.campos.to(dev), prefiltered=False, debug=debug)


class OursResult:
    pass


def run_ours(sc, cam, bg, D, grad_out=None, precomp_cov=None, precomp_col=None, scale_modifier=1.0, dev="cuda",
             debug=False, export=True):
    """Through the public Python API (GaussianRasterizer), like gaussian_renderer.render does."""
    import diff_gaussian_rasterization as dgr
    rs = settings(dgr, cam, bg, D, dev, scale_modifier, debug)
    leaf = {k: v.to(dev).clone().requires_grad_(grad_out is not None) for k, v in sc.items()}
    means2D = torch.zeros_like(leaf["means3D"], requires_grad=grad_out is not None)
    kw = {}
    extra = {}
    if precomp_col is not None:
        extra["colors"] = torch.tensor(precomp_col, device=dev, requires_grad=grad_out is not None)
        kw["colors_precomp"] = extra["colors"]
    else:
        kw["shs"] = leaf["shs"]
    if precomp_cov is not None:
        extra["cov3D"] = torch.tensor(precomp_cov, device=dev, requires_grad=grad_out is not None)
        kw["cov3D_precomp"] = extra["cov3D"]
    else:
        kw["scales"] = leaf["scales"]; kw["rotations"] = leaf["rotations"]
    r = OursResult()
    # call the autograd Function directly when we want the saved blobs for the integer contract
    color, radii = dgr.GaussianRasterizer(rs)(means3D=leaf["means3D"], means2D=means2D, opacities=leaf["opacities"], **kw)
    r.color, r.radii = color.detach().cpu(), radii.cpu()
    r.grads = None
    if grad_out is not None:
        color.backward(grad_out.to(dev))
        g = dict(means3D=leaf["means3D"].grad, means2D=means2D.grad, opacities=leaf["opacities"].grad)
        if precomp_col is None: g["sh"] = leaf["shs"].grad
        else: g["colors"] = extra["colors"].grad
        if precomp_cov is None: g["scales"] = leaf["scales"].grad; g["rotations"] = leaf["rotations"].grad
        else: g["cov3D"] = extra["cov3D"].grad
        r.grads = {k: (None if v is None else v.detach().cpu().numpy()) for k, v in g.items()}
    return r


def raw_forward(C, sc, cam, bg, D, dev="cuda", scale_modifier=1.0, precomp_cov=None, precomp_col=None, debug=False):
    """Calls a `_C`-style module (ours or the reference's) directly; returns the raw 6-tuple."""
    e = torch.Tensor([])
    t = lambda a: e if a is None else torch.as_tensor(a).to(dev)
    args = (bg.to(dev), sc["means3D"].to(dev), t(precomp_col), sc["opacities"].to(dev),
            e if precomp_cov is not None else sc["scales"].to(dev), e if precomp_cov is not None else sc["rotations"].to(dev),
            scale_modifier, t(precomp_cov), cam.viewmatrix.to(dev), cam.projmatrix.to(dev), cam.tanfovx, cam.tanfovy,
            cam.H, cam.W, e if precomp_col is not None else sc["shs"].to(dev), D, cam.campos.to(dev), False, debug)
    return args, C.rasterize_gaussians(*args)


def raw_backward(C, args, fwd, grad_out, dev="cuda"):
    (bg, means3D, colors, opac, scales, rots, mod, cov, vm, pm, tx, ty, H, W, sh, D, campos, pref, debug) = args
    R, color, radii, geom, binning, img = fwd
    return C.rasterize_gaussians_backward(bg, means3D, radii, colors, scales, rots, mod, cov, vm, pm, tx, ty,
                                          grad_out.to(dev), sh, D, campos, geom, R, binning, img, debug)


def load_reference():
    """The unmodified reference extension (oracle/_ref/ref_dgr_C*.so), or None when it was not built."""
    if not os.path.isdir(REF_DIR) or not any(f.startswith("ref_dgr_C") for f in os.listdir(REF_DIR)):
        return None, None
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    try:
        import ref_dgr_C
        import ref_knn_C
        return ref_dgr_C, ref_knn_C
    except Exception as ex:  # pragma: no cover
        print("reference extension present but not loadable:", ex)
        return None, None


def _al(x, a=128):
    return (x + a - 1) // a * a


def parse_reference_blobs(binning: torch.Tensor, img: torch.Tensor, R: int, W: int, H: int):
    """sorted values / keys from BinningState, n_contrib / ranges / final_T from ImageState."""
    N = W * H
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    b = binning.cpu().numpy()
    base = binning.data_ptr()
    o0 = _al(base) - base
    o1 = _al(base + o0 + 4 * R) - base
    o2 = _al(base + o1 + 4 * R) - base
    vals = b[o0:o0 + 4 * R].view(np.uint32).copy()
    keys = b[o2:o2 + 8 * R].view(np.uint64).copy()
    i = img.cpu().numpy()
    ib = img.data_ptr()
    p0 = _al(ib) - ib
    p1 = _al(ib + p0 + 4 * N) - ib
    p2 = _al(ib + p1 + 4 * N) - ib
    final_T = i[p0:p0 + 4 * N].view(np.float32).copy()
    n_contrib = i[p1:p1 + 4 * N].view(np.uint32).copy()
    ranges = i[p2:p2 + 8 * tiles].view(np.uint32).reshape(tiles, 2).copy()
    return dict(keys=keys, vals=vals, ranges=ranges, n_contrib=n_contrib, final_T=final_T)


def export_ours(C, fwd, P, W, H):
    R, color, radii, geom, binning, img = fwd
    keys, vals, ranges, ncontrib, finalT = C.export_binning(geom, binning, img, P, R, W, H)
    return dict(keys=keys.cpu().numpy().view(np.uint64), vals=vals.cpu().numpy().view(np.uint32),
                ranges=ranges.cpu().numpy().view(np.uint32), n_contrib=ncontrib.cpu().numpy().view(np.uint32).reshape(-1),
                final_T=finalT.cpu().numpy().reshape(-1))


def rel_err(a, b):
    """max |a-b| / max|b| : the 'relative' gradient tolerance of BASELINE.json (atomic order differs)."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    d = np.abs(a - b).max() if a.size else 0.0
    s = np.abs(b).max() if b.size else 0.0
    return d / s if s > 0 else d
