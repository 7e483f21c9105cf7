"""PLY wire format of a Gaussian model without `plyfile`.

Mirror of GaussianModel.save_ply / load_ply (reference scene/gaussian_model.py:187-266): one `vertex` element, binary
little-endian, float32 properties
    x y z nx ny nz f_dc_0..2 f_rest_0..(3K-1) opacity scale_0..2 rot_0..3          (K = (max_sh_degree+1)^2 - 1)
where the SH columns are CHANNEL-major in the file (f_rest_{c*K + k}) and coefficient-major in memory
(`_features_rest[P, K, 3]`, :213 / :245-259).  Values are stored raw (pre-activation), exactly as the reference does.

The header is parsed / written here; the payload moves as one block (np.fromfile -> pinned -> H2D) and the column
shuffles run on the GPU (csrc/io.cu gsr_gather_columns) instead of 62 strided numpy column copies.
Files with extra / reordered / non-float32 scalar properties (other exporters) are read through a numpy structured view and
converted to float32 first.  CUDA only for the tensors; no CPU fallback for the shuffles.
"""
from __future__ import annotations

import os
from typing import Dict, List, Tuple

import numpy as np
import torch

from . import load_binding

_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2", "ushort": "u2",
              "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4", "float": "f4", "float32": "f4",
              "double": "f8", "float64": "f8"}


def attribute_names(n_rest: int) -> List[str]:
    """construct_list_of_attributes (gaussian_model.py:187-199) for K = n_rest / 3 extra SH coefficients"""
    names = ["x", "y", "z", "nx", "ny", "nz"] + ["f_dc_%d" % i for i in range(3)] + ["f_rest_%d" % i for i in range(n_rest)]
    return names + ["opacity"] + ["scale_%d" % i for i in range(3)] + ["rot_%d" % i for i in range(4)]


def format_header(n: int, names: List[str]) -> bytes:
    lines = ["ply", "format binary_little_endian 1.0", "element vertex %d" % n]
    lines += ["property float %s" % a for a in names]
    return ("\n".join(lines) + "\nend_header\n").encode("ascii")


def parse_header(buf: bytes) -> Tuple[int, List[Tuple[str, str]], int, str]:
    """returns (vertex count, [(property name, numpy dtype code)], payload offset, format) for the first element"""
    end = buf.find(b"end_header")
    if not buf.startswith(b"ply") or end < 0:
        raise ValueError("not a PLY file")
    nl = buf.find(b"\n", end)
    if nl < 0:
        raise ValueError("truncated PLY header")
    fmt, count, props, in_vertex, seen = None, None, [], False, 0
    for line in buf[:end].decode("ascii", "replace").splitlines():
        tok = line.split()
        if not tok or tok[0] == "comment" or tok[0] == "obj_info":
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            seen += 1
            in_vertex = seen == 1
            if in_vertex:
                if tok[1] != "vertex":
                    raise ValueError("first PLY element is %r, expected 'vertex'" % tok[1])
                count = int(tok[2])
        elif tok[0] == "property" and in_vertex:
            if tok[1] == "list":
                raise ValueError("list properties are not part of the Gaussian model format")
            if tok[1] not in _PLY_TYPES:
                raise ValueError("unknown PLY property type %r" % tok[1])
            props.append((tok[2], _PLY_TYPES[tok[1]]))
    if fmt is None or count is None:
        raise ValueError("PLY header lacks format / vertex element")
    return count, props, nl + 1, fmt


def _index_sorted(names: List[str], prefix: str) -> List[int]:
    """columns whose name starts with prefix, ordered by the integer after the last '_' (gaussian_model.py:235-236,246-253)"""
    sel = [(int(n.split("_")[-1]), i) for i, n in enumerate(names) if n.startswith(prefix)]
    return [i for _, i in sorted(sel)]


def column_maps(names: List[str], max_sh_degree: int) -> Dict[str, List[int]]:
    """for each parameter tensor, the file column of every element in MEMORY order"""
    col = {n: i for i, n in enumerate(names)}
    K = (max_sh_degree + 1) ** 2 - 1
    rest = _index_sorted(names, "f_rest_")
    if len(rest) != 3 * K:                                              # the reference's assert, gaussian_model.py:237
        raise ValueError("PLY holds %d f_rest columns, max_sh_degree %d needs %d" % (len(rest), max_sh_degree, 3 * K))
    try:
        return {"xyz": [col["x"], col["y"], col["z"]],
                "features_dc": [col["f_dc_0"], col["f_dc_1"], col["f_dc_2"]],
                "features_rest": [rest[c * K + k] for k in range(K) for c in range(3)],    # [K,3] memory <- channel-major file
                "opacity": [col["opacity"]],
                "scaling": _index_sorted(names, "scale_"),
                "rotation": _index_sorted(names, "rot")}
    except KeyError as e:
        raise ValueError("PLY lacks property %s" % e) from None


def read_payload(path: str) -> Tuple[np.ndarray, List[str]]:
    """the vertex element as a float32 [P, n_props] array (zero-copy view of the file bytes when all properties are float32)"""
    with open(path, "rb") as f:
        head = f.read(1 << 16)
    n, props, off, fmt = parse_header(head)
    names = [p for p, _ in props]
    if fmt == "binary_little_endian":
        dt = np.dtype([(p, "<" + t) for p, t in props])
        raw = np.fromfile(path, dtype=dt, count=n, offset=off)
        if raw.shape[0] != n:
            raise ValueError("PLY payload shorter than its header says")
        if all(t == "f4" for _, t in props):
            return raw.view("<f4").reshape(n, len(props)), names
        return np.stack([raw[p].astype(np.float32) for p in names], axis=1), names
    if fmt == "ascii":
        with open(path, "rb") as f:
            f.seek(off)
            data = np.loadtxt(f, dtype=np.float64, max_rows=n, ndmin=2)
        return data[:, :len(props)].astype(np.float32), names
    raise ValueError("PLY format %r not supported (the reference writes binary_little_endian)" % fmt)


def load_ply(path: str, max_sh_degree: int = 3, device="cuda") -> Dict[str, torch.Tensor]:
    """GaussianModel.load_ply: returns the raw parameter tensors on `device`
    (xyz [P,3], features_dc [P,1,3], features_rest [P,K,3], opacity [P,1], scaling [P,3], rotation [P,4])."""
    payload, names = read_payload(path)
    maps = column_maps(names, max_sh_degree)
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("gsr_b200.ply.load_ply: CUDA device only (no CPU fallback)")
    host = torch.from_numpy(np.ascontiguousarray(payload))
    rec = host.pin_memory().to(dev, non_blocking=True) if host.numel() else host.to(dev)
    b = load_binding()
    P = rec.shape[0]
    K = (max_sh_degree + 1) ** 2 - 1
    out = {k: b.gather_columns(rec, v) for k, v in maps.items()}
    out["features_dc"] = out["features_dc"].view(P, 1, 3)
    out["features_rest"] = out["features_rest"].view(P, K, 3)
    return out


def save_ply(path: str, xyz, features_dc, features_rest, opacity, scaling, rotation) -> None:
    """GaussianModel.save_ply: writes the file the reference (and its viewers) read back"""
    P = xyz.shape[0]
    K = features_rest.shape[1]
    if not xyz.is_cuda:
        raise RuntimeError("gsr_b200.ply.save_ply: CUDA tensors only (no CPU fallback)")
    block = torch.cat([t.detach().reshape(P, t.numel() // max(P, 1)).float() for t in (xyz, features_dc, features_rest, opacity, scaling, rotation)],
                      dim=1).contiguous()                                                   # [P, 3 + 3 + 3K + 1 + 3 + 4]
    o_dc, o_rest, o_op = 3, 6, 6 + 3 * K
    cols = [0, 1, 2, -1, -1, -1] + [o_dc + c for c in range(3)] + [o_rest + k * 3 + c for c in range(3) for k in range(K)]
    cols += [o_op] + [o_op + 1 + i for i in range(3)] + [o_op + 4 + i for i in range(4)]
    b = load_binding()
    parts = [b.gather_columns(block, cols[i:i + 64]) for i in range(0, len(cols), 64)]
    rec = (parts[0] if len(parts) == 1 else torch.cat(parts, dim=1)).cpu().numpy()
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    with open(path, "wb") as f:
        f.write(format_header(P, attribute_names(3 * K)))
        f.write(np.ascontiguousarray(rec, dtype="<f4").tobytes())
