"""CPU restatement (numpy) of GaussianModel.load_ply / save_ply -- TEST INFRASTRUCTURE ONLY.

Follows scene/gaussian_model.py:201-218 (save: columns x y z nx ny nz f_dc f_rest opacity scale rot, SH flattened from the
transposed [P,3,K] view, normals zero) and :224-266 (load: columns by name, f_rest / scale / rot sorted by their numeric
suffix, f_rest reshaped (P, 3, K) then transposed to [P,K,3]).  The reference reads and writes through the `plyfile`
package (absent here and un-vendored: PARITY UNPINNED against plyfile itself); what is restated is the published PLY
format it emits for a structured float32 array: 'ply / format binary_little_endian 1.0 / element vertex N /
property float <name>... / end_header' followed by N packed little-endian records.
"""
from __future__ import annotations

import numpy as np


def _names(K3):
    l = ["x", "y", "z", "nx", "ny", "nz"] + ["f_dc_%d" % i for i in range(3)] + ["f_rest_%d" % i for i in range(K3)]
    return l + ["opacity"] + ["scale_%d" % i for i in range(3)] + ["rot_%d" % i for i in range(4)]


def save_ply(path, xyz, features_dc, features_rest, opacity, scaling, rotation):
    P = xyz.shape[0]
    f_dc = features_dc.transpose(0, 2, 1).reshape(P, 3 * features_dc.shape[1])
    f_rest = features_rest.transpose(0, 2, 1).reshape(P, 3 * features_rest.shape[1])
    attrs = np.concatenate((xyz, np.zeros_like(xyz), f_dc, f_rest, opacity, scaling, rotation), axis=1).astype("<f4")
    names = _names(f_rest.shape[1])
    assert attrs.shape[1] == len(names)
    with open(path, "wb") as f:
        f.write(("ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P).encode())
        f.write("".join("property float %s\n" % n for n in names).encode())
        f.write(b"end_header\n")
        f.write(attrs.tobytes())


def load_ply(path, max_sh_degree=3):
    with open(path, "rb") as f:
        data = f.read()
    end = data.index(b"end_header\n") + len(b"end_header\n")
    names, n = [], None
    for line in data[:end].decode().splitlines():
        t = line.split()
        if t[:1] == ["element"]:
            n = int(t[2])
        if t[:1] == ["property"]:
            assert t[1] in ("float", "float32")
            names.append(t[2])
    el = np.frombuffer(data, dtype=np.dtype([(k, "<f4") for k in names]), count=n, offset=end)
    xyz = np.stack((el["x"], el["y"], el["z"]), axis=1)
    opac = np.asarray(el["opacity"])[..., None]
    fdc = np.stack((el["f_dc_0"], el["f_dc_1"], el["f_dc_2"]), axis=1)[:, :, None]                  # (P, 3, 1)
    by_suffix = lambda pre: sorted([k for k in names if k.startswith(pre)], key=lambda x: int(x.split("_")[-1]))
    extra = by_suffix("f_rest_")
    K = (max_sh_degree + 1) ** 2 - 1
    assert len(extra) == 3 * K
    fx = np.stack([el[k] for k in extra], axis=1).reshape(n, 3, K) if extra else np.zeros((n, 3, 0), np.float32)
    scales = np.stack([el[k] for k in by_suffix("scale_")], axis=1)
    rots = np.stack([el[k] for k in by_suffix("rot")], axis=1)
    return {"xyz": xyz, "features_dc": np.ascontiguousarray(fdc.transpose(0, 2, 1)),
            "features_rest": np.ascontiguousarray(fx.transpose(0, 2, 1)), "opacity": opac, "scaling": scales, "rotation": rots}
