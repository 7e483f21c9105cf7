"""Generates tests/golden/optim_adam.npz with torch.optim.Adam ITSELF (CPU), configured as the reference configures it
(scene/gaussian_model.py:164-175: six groups, lr per group from arguments/__init__.py:79-86, lr=0.0 default, eps=1e-15),
stepped three times with seeded gradients the way train.py:138-140 does (step, zero_grad(set_to_none=True)).
Stores initial parameters, the gradients of every step and the parameters / moments after the last step.
"""
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
GROUPS = [("xyz", (3,), 0.00016 * 5.0), ("f_dc", (1, 3), 0.0025), ("f_rest", (15, 3), 0.0025 / 20.0), ("opacity", (1,), 0.05),
          ("scaling", (3,), 0.005), ("rotation", (4,), 0.001)]
P, STEPS = 37, 3


def main():
    g = torch.Generator().manual_seed(3)
    params = {n: torch.nn.Parameter(torch.randn(P, *s, generator=g)) for n, s, _ in GROUPS}
    out = {"names": np.array([n for n, _, _ in GROUPS]), "lrs": np.array([lr for _, _, lr in GROUPS], np.float64),
           "steps": np.int32(STEPS)}
    for n in params:
        out["p0_" + n] = params[n].detach().numpy().copy()
    opt = torch.optim.Adam([{"params": [params[n]], "lr": lr, "name": n} for n, _, lr in GROUPS], lr=0.0, eps=1e-15)
    for t in range(STEPS):
        for n in params:
            gr = torch.randn(params[n].shape, generator=g) * (10.0 ** float(torch.randint(-6, 0, (1,), generator=g)))
            if t == 1:
                gr[::5] = 0.0                                   # culled Gaussians: zero gradient rows
            params[n].grad = gr
            out["g%d_%s" % (t, n)] = gr.numpy().copy()
        opt.step()
        opt.zero_grad(set_to_none=True)
    for n in params:
        st = opt.state[params[n]]
        out["p_" + n] = params[n].detach().numpy()
        out["m_" + n] = st["exp_avg"].numpy()
        out["v_" + n] = st["exp_avg_sq"].numpy()
    p = os.path.join(HERE, "optim_adam.npz")
    np.savez_compressed(p, **out)
    print("wrote", p)


if __name__ == "__main__":
    main()
