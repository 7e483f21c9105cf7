"""CPU restatement (numpy float32) of the optimizer step and the densification statistics -- TEST INFRASTRUCTURE ONLY.

adam_step follows what the reference's `gaussians.optimizer.step()` executes (train.py:138-140): torch.optim.Adam as built at
scene/gaussian_model.py:164-175 (lr per group, betas (0.9, 0.999), eps 1e-15, no weight decay, no amsgrad).  torch is a pinned
third-party dependency (environment.yml: pytorch 1.12.1; here 2.11); its published single-tensor algorithm
(torch/optim/adam.py `_single_tensor_adam`) is restated:
    step += 1;  m = lerp(m, g, 1 - b1);  v = v b2 + (1 - b2) g g
    p += -(lr / (1 - b1^step)) * m / (sqrt(v) / sqrt(1 - b2^step) + eps)      (bias corrections in Python doubles)
Pinned by tests/golden/optim_adam.npz, produced by torch.optim.Adam itself on the CPU (tests/golden/make_optim_golden.py).
densify_stats follows train.py:131-132 and scene/gaussian_model.py:415-417.
"""
from __future__ import annotations

import math

import numpy as np

F = np.float32


def adam_step(p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-15, visible=None):
    """one update of one group, in place on float32 arrays; `step` is the 1-based count after this update"""
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    neg_step = F(-(lr / bc1))
    bc2s = F(math.sqrt(bc2))
    m2 = (m + F(1.0 - beta1) * (g - m)).astype(F)
    v2 = (v * F(beta2) + F(1.0 - beta2) * g * g).astype(F)
    denom = (np.sqrt(v2) / bc2s + F(eps)).astype(F)
    p2 = (p + neg_step * (m2 / denom)).astype(F)
    if visible is not None:
        rows = p.size // visible.size
        keep = np.repeat(visible.astype(bool), rows).reshape(p.shape)
        m2, v2, p2 = np.where(keep, m2, m), np.where(keep, v2, v), np.where(keep, p2, p)
    p[...] = p2; m[...] = m2; v[...] = v2


def densify_stats(radii, dmean2D, max_radii2D, grad_accum, denom):
    vis = radii > 0
    max_radii2D[vis] = np.maximum(max_radii2D[vis], radii[vis].astype(F))
    grad_accum[vis] += np.sqrt((dmean2D[vis, :2].astype(F) ** 2).sum(axis=1, dtype=F))
    denom[vis] += F(1)
