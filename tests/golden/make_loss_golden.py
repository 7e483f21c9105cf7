"""Generates tests/golden/loss_l1_ssim.npz by IMPORTING the reference's own loss functions from /root/reference
(run in the build container on the CPU; the reference is not on the GPU box):
  utils/loss_utils.py:17-18 l1_loss, :38-63 ssim   combined as train.py:104-105 with lambda_dssim = 0.2.
Stores two small image pairs, the reference's l1 / ssim / loss values and torch autograd's dloss/dimage.
"""
import os
import sys

import numpy as np
import torch

REF = os.environ.get("REF", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [REF]
from utils.loss_utils import l1_loss, ssim  # noqa: E402

LAMBDA = 0.2


def main():
    out = {"lambda_dssim": np.float32(LAMBDA)}
    g = torch.Generator().manual_seed(11)
    for name, (C, H, W) in {"a": (3, 37, 45), "b": (3, 64, 96), "c": (1, 9, 7)}.items():
        x = torch.rand(C, H, W, generator=g)
        t = (x + 0.25 * torch.randn(C, H, W, generator=g)).clamp(0, 1)      # correlated target: SSIM away from 0
        x.requires_grad_(True)
        l1 = l1_loss(x, t)
        s = ssim(x, t)
        loss = (1.0 - LAMBDA) * l1 + LAMBDA * (1.0 - s)
        loss.backward()
        out[name + "_image"] = x.detach().numpy()
        out[name + "_target"] = t.numpy()
        out[name + "_l1"] = np.float32(l1.item())
        out[name + "_ssim"] = np.float32(s.item())
        out[name + "_loss"] = np.float32(loss.item())
        out[name + "_grad"] = x.grad.numpy()
    p = os.path.join(HERE, "loss_l1_ssim.npz")
    np.savez_compressed(p, **out)
    print("wrote", p, {k: (v.shape if hasattr(v, "shape") and v.ndim else float(v)) for k, v in out.items()})


if __name__ == "__main__":
    main()
