"""Independent fp64 torch-autograd restatement of the rasterizer forward -- TEST INFRASTRUCTURE ONLY.

Used to cross-check the hand-derived backward of ``gsr_oracle.c`` (which restates
backward.cu literally) against automatic differentiation of the restated forward
(SURVEY.md section 8c / Appendix A):
  (i)   min(0.99, .) is straight-through,
  (ii)  skip / stop masks and tile-rect membership are constants,
  (iii) means2D is an additive NDC offset (its grad carries the 0.5*W, 0.5*H factors),
  (iv)  the frustum-clamped t.x / t.y are detached (backward.cu:262-264),
  (v)   the quaternion is used un-normalised.
The integer work (radii, tile lists) is taken from the C oracle's Forward state so both
oracles composite exactly the same lists.  Small scenes only (dense per-tile tensors).
"""
from __future__ import annotations

import numpy as np
import torch

from . import TILE

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435]


def _sh_rgb(D, sh, dirs):
    x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
    res = C0 * sh[:, 0]
    if D > 0:
        res = res - C1 * y * sh[:, 1] + C1 * z * sh[:, 2] - C1 * x * sh[:, 3]
        if D > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            res = (res + C2[0] * xy * sh[:, 4] + C2[1] * yz * sh[:, 5] + C2[2] * (2 * zz - xx - yy) * sh[:, 6]
                   + C2[3] * xz * sh[:, 7] + C2[4] * (xx - yy) * sh[:, 8])
            if D > 2:
                res = (res + C3[0] * y * (3 * xx - yy) * sh[:, 9] + C3[1] * xy * z * sh[:, 10]
                       + C3[2] * y * (4 * zz - xx - yy) * sh[:, 11] + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[:, 12]
                       + C3[4] * x * (4 * zz - xx - yy) * sh[:, 13] + C3[5] * z * (xx - yy) * sh[:, 14]
                       + C3[6] * x * (xx - 3 * yy) * sh[:, 15])
    return torch.clamp_min(res + 0.5, 0.0)


def render_autograd(st, dL_dpix):
    """st: oracle.Forward (C oracle state).  Returns (color[3,H,W] fp64, grads dict fp64)."""
    i = st.inputs
    dt = torch.float64
    T_ = lambda a: None if a is None else torch.tensor(np.asarray(a), dtype=dt)
    means = T_(i["means3D"]).requires_grad_(True)
    P = means.shape[0]
    m2d = torch.zeros(P, 2, dtype=dt, requires_grad=True)
    opac = torch.tensor(st.conic_o[:, 3].copy(), dtype=dt)
    # opacity of culled gaussians is not recorded in conic_o; they never contribute anyway
    opac.requires_grad_(True)
    V = T_(i["V"]).reshape(4, 4)      # V[c, r] = element (r, c)
    PR = T_(i["PR"]).reshape(4, 4)
    cam = T_(i["cam"])
    W, H = st.W, st.H
    tanx, tany = i["tanfovx"], i["tanfovy"]
    fx, fy = W / (2.0 * tanx), H / (2.0 * tany)
    leaves = dict(means3D=means, means2D=m2d, opacities=opac)

    ones = torch.ones(P, 1, dtype=dt)
    mh = torch.cat([means, ones], 1)
    pv = mh @ V                         # row-vector convention
    ph = mh @ PR
    pw = 1.0 / (ph[:, 3] + 1e-7)
    ndc = ph[:, :2] * pw[:, None] + m2d
    pix = torch.stack([((ndc[:, 0] + 1.0) * W - 1.0) * 0.5, ((ndc[:, 1] + 1.0) * H - 1.0) * 0.5], 1)

    if i["cov3D_precomp"] is not None:
        c6 = T_(i["cov3D_precomp"]).requires_grad_(True)
        leaves["cov3D"] = c6
        Sig = torch.stack([c6[:, 0], c6[:, 1], c6[:, 2], c6[:, 1], c6[:, 3], c6[:, 4], c6[:, 2], c6[:, 4], c6[:, 5]],
                          1).reshape(P, 3, 3)
    else:
        s = T_(i["scales"]).requires_grad_(True)
        q = T_(i["rotations"]).requires_grad_(True)
        leaves["scales"], leaves["rotations"] = s, q
        r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        Rq = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                          2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                          2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).reshape(P, 3, 3)
        sm = s * i["scale_modifier"]
        Sig = Rq @ torch.diag_embed(sm * sm) @ Rq.transpose(1, 2)

    tz = pv[:, 2]
    limx, limy = 1.3 * tanx, 1.3 * tany
    rx, ry = pv[:, 0] / tz, pv[:, 1] / tz
    tx = torch.where((rx >= -limx) & (rx <= limx), pv[:, 0], (rx.clamp(-limx, limx) * tz).detach())
    ty = torch.where((ry >= -limy) & (ry <= limy), pv[:, 1], (ry.clamp(-limy, limy) * tz).detach())
    zero = torch.zeros_like(tz)
    J = torch.stack([fx / tz, zero, -fx * tx / (tz * tz), zero, fy / tz, -fy * ty / (tz * tz)], 1).reshape(P, 2, 3)
    Rw = V[:3, :3].t()                  # W2C rotation, (row, col)
    A = J @ Rw
    cov = A @ Sig @ A.transpose(1, 2)
    a, b, c = cov[:, 0, 0] + 0.3, cov[:, 0, 1], cov[:, 1, 1] + 0.3
    det = a * c - b * b
    conic = torch.stack([c / det, -b / det, a / det], 1)

    if i["colors_precomp"] is not None:
        rgb = T_(i["colors_precomp"]).requires_grad_(True)
        leaves["colors"] = rgb
    else:
        sh = T_(i["shs"]).requires_grad_(True)
        leaves["sh"] = sh
        d = means - cam
        rgb = _sh_rgb(st.D, sh, d / d.norm(dim=1, keepdim=True))

    bg = T_(i["bg"])
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    out = torch.zeros(3, H, W, dtype=dt)
    ranges = st.ranges.astype(np.int64)
    lst = torch.tensor(st.vals.astype(np.int64))
    pieces = []
    for t in range(gx * gy):
        x0, y0 = (t % gx) * TILE, (t // gx) * TILE
        xs = torch.arange(x0, min(x0 + TILE, W)); ys = torch.arange(y0, min(y0 + TILE, H))
        yy, xx = torch.meshgrid(ys, xs, indexing="ij")
        pxf, pyf = xx.reshape(-1).to(dt), yy.reshape(-1).to(dt)
        ids = lst[ranges[t, 0]:ranges[t, 1]]
        n = ids.numel()
        if n == 0:
            col = bg[:, None].expand(3, pxf.numel())
        else:
            dx = pix[ids, 0][None, :] - pxf[:, None]
            dy = pix[ids, 1][None, :] - pyf[:, None]
            cn = conic[ids]
            power = -0.5 * (cn[None, :, 0] * dx * dx + cn[None, :, 2] * dy * dy) - cn[None, :, 1] * dx * dy
            araw = opac[ids][None, :] * torch.exp(power)
            alpha = araw + (araw.clamp(max=0.99) - araw).detach()
            keep = ((power <= 0) & (alpha >= 1.0 / 255.0)).detach()
            aeff = torch.where(keep, alpha, torch.zeros_like(alpha))
            one_m = 1.0 - aeff
            Tincl = torch.cumprod(one_m, 1)
            Texcl = torch.cat([torch.ones_like(Tincl[:, :1]), Tincl[:, :-1]], 1)
            stop = (keep & (Tincl < 1e-4)).detach()
            alive = (torch.cumsum(stop.to(torch.int64), 1) == 0)
            w = torch.where(alive & keep, aeff * Texcl, torch.zeros_like(aeff))
            Tfin = torch.where(alive, Tincl, torch.zeros_like(Tincl))
            # final T = T after the last blended entry = product over alive&keep entries
            Tlast = torch.prod(torch.where(alive, one_m, torch.ones_like(one_m)), 1)
            col = (w @ rgb[ids]).t() + Tlast[None, :] * bg[:, None]
            del Tfin
        pieces.append((yy.reshape(-1), xx.reshape(-1), col))
    ys = torch.cat([p[0] for p in pieces]); xs = torch.cat([p[1] for p in pieces])
    cols = torch.cat([p[2] for p in pieces], 1)
    out = torch.zeros(3, H, W, dtype=dt).index_put((torch.arange(3)[:, None].expand(3, ys.numel()),
                                                   ys[None, :].expand(3, -1), xs[None, :].expand(3, -1)), cols)
    loss = (out * torch.tensor(np.asarray(dL_dpix), dtype=dt).reshape(3, H, W)).sum()
    names = list(leaves)
    gs = torch.autograd.grad(loss, [leaves[k] for k in names], allow_unused=True)
    grads = {k: (None if g is None else g.detach().numpy()) for k, g in zip(names, gs)}
    return out.detach().numpy(), grads
