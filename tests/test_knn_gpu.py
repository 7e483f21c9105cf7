"""distCUDA2 parity (GPU): ours vs the brute-force oracle, the committed reference output
(tests/golden/ref_knn20k.npz, produced by the unmodified simple-knn on a B200) and, when built, the
reference extension itself.  The result is an exact 3-NN, so the bar is bit-exact; the CPU oracle
does not fuse multiply-adds, so against it the bar is 2 ulp."""
import os

import numpy as np
import pytest
import torch

import util
from gsr_b200 import synth

pytestmark = pytest.mark.gpu


def _dist2(pts):
    from simple_knn._C import distCUDA2
    return distCUDA2(pts.cuda()).cpu().numpy()


@pytest.mark.parametrize("P", [4, 5, 257, 5000])
def test_vs_bruteforce_oracle(P):
    import oracle
    g = torch.Generator().manual_seed(P)
    pts = torch.rand(P, 3, generator=g) * 2 - 1
    if P >= 257:
        pts[7] = pts[3]                      # coincident points count as distance 0
        pts[100:110, 2] = 0.25               # a coplanar cluster
    ours = _dist2(pts)
    ref = oracle.dist2(pts.numpy(), "brute")
    assert np.allclose(ours, ref, rtol=3e-7, atol=0)


def test_golden_reference_20k():
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_knn20k.npz"))
    pts = synth.make_scene(20000, 0, 1.0, seed=43)["means3D"]
    ours = _dist2(pts)
    assert np.array_equal(ours, g["dist2"]), "max rel diff %.3e" % np.abs(ours / g["dist2"] - 1).max()


def test_vs_reference_extension_300k():
    _, knn = util.load_reference()
    if knn is None:
        pytest.skip("oracle/_ref not built on this machine")
    pts = synth.make_scene(300000, 0, 1.5, seed=47)["means3D"].cuda()
    from simple_knn._C import distCUDA2
    assert torch.equal(distCUDA2(pts), knn.distCUDA2(pts))


def test_negative_only_and_offset_cloud():
    """bounding box always contains the origin (reduction init 0, simple_knn.cu:191-199)"""
    import oracle
    g = torch.Generator().manual_seed(5)
    pts = torch.rand(3000, 3, generator=g) * 0.5 + 3.0
    assert np.allclose(_dist2(pts), oracle.dist2(pts.numpy(), "brute"), rtol=3e-7, atol=0)
    assert np.allclose(_dist2(-pts), oracle.dist2((-pts).numpy(), "brute"), rtol=3e-7, atol=0)
