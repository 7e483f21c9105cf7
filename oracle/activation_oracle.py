"""CPU restatement (numpy float32) of the model's parameter activations and their backward -- TEST INFRASTRUCTURE ONLY.

Follows GaussianModel.setup_functions / get_scaling / get_rotation / get_opacity (scene/gaussian_model.py:33-40,100-125):
torch.exp, torch.nn.functional.normalize (x / max(|x|_2, 1e-12) along dim 1), torch.sigmoid.  Pinned by
tests/golden/activations.npz, produced with those torch functions and torch autograd (tests/golden/make_activation_golden.py).
"""
import numpy as np

F = np.float32
EPS = F(1e-12)


def forward(scaling, rotation, opacity):
    n = np.maximum(np.sqrt((rotation * rotation).sum(axis=1, keepdims=True, dtype=F)), EPS)
    return np.exp(scaling).astype(F), (rotation / n).astype(F), (F(1) / (F(1) + np.exp(-opacity))).astype(F)


def backward(scaling, rotation, opacity, g_s, g_q, g_o):
    s, q, o = forward(scaling, rotation, opacity)
    nr = np.sqrt((rotation * rotation).sum(axis=1, keepdims=True, dtype=F))
    n = np.maximum(nr, EPS)
    dot = np.where(nr >= EPS, (q * g_q).sum(axis=1, keepdims=True, dtype=F), F(0))
    return (g_s * s).astype(F), ((g_q - q * dot) / n).astype(F), (g_o * (F(1) - o) * o).astype(F)
