"""Generates tests/golden/activations.npz with the torch functions the reference binds as its activations
(scene/gaussian_model.py:33-40: torch.exp, torch.sigmoid, torch.nn.functional.normalize) and torch autograd, on the CPU."""
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    g = torch.Generator().manual_seed(8)
    P = 301
    s = (torch.randn(P, 3, generator=g) * 1.5 - 4.0).requires_grad_(True)
    q = torch.randn(P, 4, generator=g)
    q[::50] *= 1e-3
    q[7] = 0.0                                            # |q| < eps: the clamp_min branch
    q.requires_grad_(True)
    o = (torch.randn(P, 1, generator=g) * 3.0).requires_grad_(True)
    ys, yq, yo = torch.exp(s), torch.nn.functional.normalize(q), torch.sigmoid(o)
    gs, gq, go = (torch.randn(t.shape, generator=g) for t in (ys, yq, yo))
    torch.autograd.backward((ys, yq, yo), (gs, gq, go))
    out = dict(scaling=s, rotation=q, opacity=o, y_scaling=ys, y_rotation=yq, y_opacity=yo, g_scaling=gs, g_rotation=gq, g_opacity=go,
               d_scaling=s.grad, d_rotation=q.grad, d_opacity=o.grad)
    p = os.path.join(HERE, "activations.npz")
    np.savez_compressed(p, **{k: v.detach().numpy() for k, v in out.items()})
    print("wrote", p)


if __name__ == "__main__":
    main()
