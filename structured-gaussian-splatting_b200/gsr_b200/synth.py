"""Synthetic scenes and cameras for parity tests and bench.py.

Everything is sampled on the CPU with a seeded ``torch.Generator`` so the bits are
identical on every box and rank, then moved to the requested device.

Camera matrices restate the reference's builders (utils/graphics_utils.py:38-71,
scene/cameras.py:54-57): row-vector convention, ``viewmatrix = W2C.T``,
``projmatrix = viewmatrix @ P.T``, ``campos = inverse(viewmatrix)[3, :3]``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np
import torch


@dataclass
class SceneSpec:
    name: str
    P: int
    D: int          # SH degree
    W: int
    H: int
    half: float     # cube half-size
    radius: float   # camera ring radius
    fovx_deg: float
    ln_scale: float # mean of log scale
    V: int = 1      # cameras on the ring
    bg_random: bool = False


# BASELINE.json configs (C2/C3 share the scene; C5 doubles P)
CONFIGS = {
    "C1": SceneSpec("C1", 10_000, 0, 256, 256, 1.0, 3.0, 50.0, math.log(0.02), 1, True),
    "C2": SceneSpec("C2", 3_000_000, 3, 1920, 1080, 1.5, 4.0, 60.0, math.log(0.005), 1),
    "C3": SceneSpec("C3", 3_000_000, 3, 1920, 1080, 1.5, 4.0, 60.0, math.log(0.005), 1),
    "C4": SceneSpec("C4", 65_536, 0, 512, 512, 1.0, 3.0, 50.0, math.log(0.012), 8),
    "C5": SceneSpec("C5", 6_000_000, 3, 1920, 1080, 1.5, 4.0, 60.0, math.log(0.005), 64),
}


def make_scene(P: int, D: int, half: float = 1.5, ln_scale: float = math.log(0.005), seed: int = 0,
               device="cpu"):
    """Returns dict(means3D[P,3], scales[P,3], rotations[P,4], opacities[P,1], shs[P,(D+1)^2,3]) fp32."""
    g = torch.Generator().manual_seed(seed)
    M = (D + 1) ** 2
    means = (torch.rand(P, 3, generator=g) * 2 - 1) * half
    scales = torch.exp(torch.randn(P, 3, generator=g) * 0.5 + ln_scale)
    rot = torch.randn(P, 4, generator=g)
    rot = rot / rot.norm(dim=1, keepdim=True)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 1.5 - 1.0)
    shs = torch.empty(P, M, 3)
    shs[:, 0, :] = torch.randn(P, 3, generator=g)
    if M > 1:
        shs[:, 1:, :] = torch.randn(P, M - 1, 3, generator=g) * 0.05
    out = dict(means3D=means, scales=scales, rotations=rot, opacities=opac, shs=shs)
    return {k: v.float().contiguous().to(device) for k, v in out.items()}


def projection_matrix(znear, zfar, fovx, fovy):
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = ty * znear, tx * znear
    bottom, left = -top, -right
    Pm = torch.zeros(4, 4)
    Pm[0, 0] = 2.0 * znear / (right - left)
    Pm[1, 1] = 2.0 * znear / (top - bottom)
    Pm[0, 2] = (right + left) / (right - left)
    Pm[1, 2] = (top + bottom) / (top - bottom)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


@dataclass
class Camera:
    W: int
    H: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor   # [4,4]
    projmatrix: torch.Tensor   # [4,4]
    campos: torch.Tensor       # [3]

    def to(self, device):
        return Camera(self.W, self.H, self.tanfovx, self.tanfovy, self.viewmatrix.to(device),
                      self.projmatrix.to(device), self.campos.to(device))


def make_camera(k: int, V: int, W: int, H: int, radius: float = 4.0, fovx_deg: float = 60.0,
                znear: float = 0.01, zfar: float = 100.0, device="cpu") -> Camera:
    """Camera k of V on a ring of `radius` in the xz-plane, looking at the origin, +y down (COLMAP)."""
    th = 2.0 * math.pi * k / V
    c = np.array([radius * math.sin(th), 0.0, radius * math.cos(th)])
    fwd = -c / np.linalg.norm(c)
    down = np.array([0.0, 1.0, 0.0])
    right = np.cross(down, fwd)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    R_c2w = np.stack([right, down, fwd], axis=1)          # columns = camera axes in world
    w2c = np.zeros((4, 4))
    w2c[:3, :3] = R_c2w.T
    w2c[:3, 3] = -R_c2w.T @ c
    w2c[3, 3] = 1.0
    view = torch.tensor(np.float32(w2c)).transpose(0, 1).contiguous()
    fovx = math.radians(fovx_deg)
    tanx = math.tan(fovx * 0.5)
    tany = tanx * H / W
    fovy = 2.0 * math.atan(tany)
    proj = projection_matrix(znear, zfar, fovx, fovy).transpose(0, 1)
    full = (view.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    campos = view.inverse()[3, :3].contiguous()
    return Camera(W, H, tanx, tany, view.to(device), full.to(device), campos.to(device))


def spec_scene(spec: SceneSpec, seed: int = 0, device="cpu", P: int | None = None):
    return make_scene(P or spec.P, spec.D, spec.half, spec.ln_scale, seed, device)


def spec_camera(spec: SceneSpec, k: int = 0, device="cpu") -> Camera:
    return make_camera(k, spec.V, spec.W, spec.H, spec.radius, spec.fovx_deg, device=device)


def spec_background(spec: SceneSpec, device="cpu"):
    if spec.bg_random:
        return torch.rand(3, generator=torch.Generator().manual_seed(1)).to(device)
    return torch.zeros(3, device=device)


def grad_image(W: int, H: int, seed: int = 2, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(3, H, W, generator=g) / (3.0 * H * W)).to(device)
