"""Times gsr_b200.ply.load_ply / save_ply against the numpy restatement of the reference's loader (oracle/ply_oracle.py, which
ends with the same .cuda() uploads GaussianModel.load_ply does) on a 1 M-Gaussian SH-3 model (248 MB file in /tmp)."""
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "structured-gaussian-splatting_b200"), ROOT]
from gsr_b200 import ply  # noqa: E402
from oracle import ply_oracle  # noqa: E402


def main():
    P = int(os.environ.get("P", 1_000_000))
    g = torch.Generator().manual_seed(0)
    m = dict(xyz=torch.randn(P, 3, generator=g), features_dc=torch.randn(P, 1, 3, generator=g),
             features_rest=torch.randn(P, 15, 3, generator=g), opacity=torch.randn(P, 1, generator=g),
             scaling=torch.randn(P, 3, generator=g), rotation=torch.randn(P, 4, generator=g))
    dev = {k: v.cuda() for k, v in m.items()}
    d = tempfile.mkdtemp()
    path = os.path.join(d, "point_cloud.ply")
    res = {"P": P, "file_bytes": 248 * P}
    ply.save_ply(path, **dev)                                   # warm-up (binding load, page cache)
    ply.load_ply(path)
    torch.cuda.synchronize()
    t = time.perf_counter(); ply.save_ply(path, **dev); res["save_ply_s"] = time.perf_counter() - t
    t = time.perf_counter(); got = ply.load_ply(path); torch.cuda.synchronize(); res["load_ply_s"] = time.perf_counter() - t
    t = time.perf_counter()
    ref = {k: torch.tensor(v, dtype=torch.float, device="cuda") for k, v in ply_oracle.load_ply(path).items()}
    torch.cuda.synchronize(); res["numpy_load_s"] = time.perf_counter() - t
    t = time.perf_counter(); ply_oracle.save_ply(path + ".ref", **{k: v.numpy() for k, v in m.items()})
    res["numpy_save_s"] = time.perf_counter() - t
    res["identical"] = all(torch.equal(got[k], ref[k]) for k in got) and open(path, "rb").read() == open(path + ".ref", "rb").read()
    os.remove(path); os.remove(path + ".ref"); os.rmdir(d)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
