"""Fused optimizer step and densification statistics of the training loop.

FusedAdam is a drop-in for the optimizer the reference builds at scene/gaussian_model.py:164-175
(`torch.optim.Adam(l, lr=0.0, eps=1e-15)` over the groups xyz / f_dc / f_rest / opacity / scaling / rotation): same
constructor, same `param_groups` (the reference rewrites `group["lr"]` per step, gaussian_model.py:179-185) and the same
per-parameter state keys `step`, `exp_avg`, `exp_avg_sq` (densification replaces and concatenates them,
gaussian_model.py:270-355), so `state_dict()` / `load_state_dict()` interchange with torch.optim.Adam.  `step()` updates
every group with ONE kernel launch (csrc/optim.cu, 28 B per element) instead of seven foreach passes.
`step(visible=mask)` is the sparse variant: only Gaussians whose flag is set are touched.

densify_stats() is train.py:131-132 + GaussianModel.add_densification_stats (gaussian_model.py:415-417) in one launch.
CUDA tensors only; there is no CPU fallback.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import load_binding


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        if lr < 0.0 or eps < 0.0 or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0):
            raise ValueError("FusedAdam: invalid hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps))

    @torch.no_grad()
    def step(self, closure=None, visible: Optional[torch.Tensor] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        binding = load_binding()
        by_hyper = {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda:
                    raise RuntimeError("FusedAdam: CUDA parameters only (no CPU fallback)")
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                rows = p.numel() // p.shape[0] if (p.dim() > 0 and p.shape[0] > 0) else 1
                key = (group["betas"][0], group["betas"][1], group["eps"])
                lists = by_hyper.setdefault(key, ([], [], [], [], [], [], []))
                for lst, v in zip(lists, (p, p.grad, st["exp_avg"], st["exp_avg_sq"], int(st["step"]), float(group["lr"]), rows)):
                    lst.append(v)
        empty = torch.empty(0)
        for (b1, b2, eps), (ps, gs, ms, vs, steps, lrs, rows) in by_hyper.items():
            binding.adam_step(ps, gs, ms, vs, steps, lrs, rows, b1, b2, eps, empty if visible is None else visible)
        return loss


def densify_stats(radii: torch.Tensor, viewspace_grad: torch.Tensor, max_radii2D: torch.Tensor,
                  xyz_gradient_accum: torch.Tensor, denom: torch.Tensor) -> None:
    """In place, for the filter radii > 0:  max_radii2D = max(max_radii2D, radii);
    xyz_gradient_accum += |viewspace_grad[:, :2]|;  denom += 1."""
    load_binding().densify_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom)
