"""Generates tests/golden/sub_sh_cov.npz by IMPORTING the reference's own torch restatements of two
kernel sub-steps from /root/reference (run in the build container; the reference is not on the GPU box):
  utils/sh_utils.py:57-112 eval_sh  + clamp_min(. + 0.5, 0)  (gaussian_renderer/__init__.py:73-78)  == computeColorFromSH
  scene/gaussian_model.py:25-29 build_covariance_from_scaling_rotation                                 == computeCov3D
The reference helpers allocate with device="cuda"; torch.zeros is wrapped to drop that argument so the
unmodified functions run on the CPU.
"""
import os
import sys

import numpy as np
import torch

REF = os.environ.get("REF", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [REF, os.path.join(ROOT, "structured-gaussian-splatting_b200")]

_zeros = torch.zeros


def _cpu_zeros(*a, **k):
    k.pop("device", None)
    return _zeros(*a, **k)


torch.zeros = _cpu_zeros
from utils.sh_utils import eval_sh  # noqa: E402
from utils.general_utils import build_scaling_rotation, strip_symmetric  # noqa: E402
from gsr_b200 import synth  # noqa: E402


def main():
    P, mod = 500, 0.8
    sc = synth.make_scene(P, 3, 1.0, np.log(0.05), seed=77)
    cam = synth.make_camera(0, 1, 64, 64, 3.0, 50.0)
    shs = sc["shs"]                                                    # [P, 16, 3]
    dirs = sc["means3D"] - cam.campos[None, :]
    dirs = dirs / dirs.norm(dim=1, keepdim=True)
    rgb = torch.clamp_min(eval_sh(3, shs.transpose(1, 2), dirs) + 0.5, 0.0)
    # build_covariance_from_scaling_rotation (scene/gaussian_model.py:25-29), rotations are unit quaternions
    L = build_scaling_rotation(mod * sc["scales"], sc["rotations"])
    cov6 = strip_symmetric(L @ L.transpose(1, 2))
    np.savez_compressed(os.path.join(HERE, "sub_sh_cov.npz"), means=sc["means3D"].numpy(), shs=shs.numpy(),
                        scales=sc["scales"].numpy(), rots=sc["rotations"].numpy(), campos=cam.campos.numpy(),
                        mod=np.float32(mod), rgb=rgb.numpy(), cov6=cov6.numpy())
    print("wrote sub_sh_cov.npz", rgb.shape, cov6.shape)


if __name__ == "__main__":
    main()
