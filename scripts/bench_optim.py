"""Times one optimizer step over a C3-sized model (3 M Gaussians, SH-3: 59 floats each) for gsr_b200.optim.FusedAdam,
torch.optim.Adam (foreach, what the reference runs) and torch.optim.Adam(fused=True); plus densify_stats."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "structured-gaussian-splatting_b200"), ROOT]
from gsr_b200.optim import FusedAdam, densify_stats  # noqa: E402

GROUPS = [("xyz", (3,), 0.0008), ("f_dc", (1, 3), 0.0025), ("f_rest", (15, 3), 0.000125), ("opacity", (1,), 0.05),
          ("scaling", (3,), 0.005), ("rotation", (4,), 0.001)]


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    P = int(os.environ.get("P", 3_000_000))
    res = {"P": P, "floats_per_gaussian": 59, "algorithmic_bytes": 28 * 59 * P}
    for label, make in (("fused_adam_ms", lambda l: FusedAdam(l, lr=0.0, eps=1e-15)),
                        ("torch_adam_foreach_ms", lambda l: torch.optim.Adam(l, lr=0.0, eps=1e-15)),
                        ("torch_adam_fused_ms", lambda l: torch.optim.Adam(l, lr=0.0, eps=1e-15, fused=True))):
        params = [torch.nn.Parameter(torch.randn(P, *s, device="cuda")) for _, s, _ in GROUPS]
        for p in params:
            p.grad = torch.randn_like(p) * 1e-4
        opt = make([{"params": [p], "lr": lr, "name": n} for p, (n, _, lr) in zip(params, GROUPS)])
        res[label] = timeit(opt.step)
        del opt, params
        torch.cuda.empty_cache()
    res["fused_adam_gbs"] = res["algorithmic_bytes"] / res["fused_adam_ms"] / 1e6
    radii = torch.randint(0, 30, (P,), dtype=torch.int32, device="cuda")
    dm = torch.randn(P, 3, device="cuda")
    mr, acc, den = torch.zeros(P, device="cuda"), torch.zeros(P, 1, device="cuda"), torch.zeros(P, 1, device="cuda")
    res["densify_stats_ms"] = timeit(lambda: densify_stats(radii, dm, mr, acc, den))

    def torch_stats():
        vis = radii > 0
        mr[vis] = torch.max(mr[vis], radii[vis].float())
        acc[vis] += torch.norm(dm[vis, :2], dim=-1, keepdim=True)
        den[vis] += 1
    res["torch_stats_ms"] = timeit(torch_stats)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
