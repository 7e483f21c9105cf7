"""Generates tests/golden/ref_*.npz: outputs of the UNMODIFIED reference CUDA extension (oracle/_ref,
built by oracle/build_ref.sh from /root/reference) on a B200, for the seeded synthetic configs.

Run on the GPU box (the reference has no CPU path):
    gpurun -- 'python tests/golden/make_ref_golden.py gpurun_out/golden'
then copy gpurun_out/golden/*.npz to tests/golden/ and commit them.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "structured-gaussian-splatting_b200"), os.path.join(ROOT, "tests")]
import util  # noqa: E402
from gsr_b200 import synth  # noqa: E402


def one(ref, name, sc, cam, bg, D, out_dir):
    a, f = util.raw_forward(ref, sc, cam, bg, D)
    e = util.parse_reference_blobs(f[4], f[5], f[0], cam.W, cam.H)
    go = synth.grad_image(cam.W, cam.H, seed=6)
    g = util.raw_backward(ref, a, f, go)
    names = ["means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations"]
    d = dict(R=np.int64(f[0]), color=f[1].cpu().numpy(), radii=f[2].cpu().numpy(), **e)
    for n, x in zip(names, g):
        d["grad_" + n] = x.cpu().numpy()
    np.savez_compressed(os.path.join(out_dir, "ref_%s.npz" % name), **d)
    print(name, "R =", f[0], "visible =", int((f[2] > 0).sum()))


def main():
    out_dir = sys.argv[1] if len(sys.argv) > 1 else HERE
    os.makedirs(out_dir, exist_ok=True)
    ref, knn = util.load_reference()
    assert ref is not None, "oracle/_ref is not built"
    spec = synth.CONFIGS["C1"]
    one(ref, "C1", synth.spec_scene(spec), synth.spec_camera(spec), synth.spec_background(spec), spec.D, out_dir)
    sc, cam, bg = util.small_scene(3000, 3, 200, 136, 41)
    one(ref, "small3", sc, cam, bg, 3, out_dir)
    pts = synth.make_scene(20000, 0, 1.0, seed=43)["means3D"]
    d2 = knn.distCUDA2(pts.cuda()).cpu().numpy()
    np.savez_compressed(os.path.join(out_dir, "ref_knn20k.npz"), dist2=d2)


if __name__ == "__main__":
    main()
