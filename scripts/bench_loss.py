"""Times the fused L1 + D-SSIM loss (value + gradient) against the plain torch formulation at 3x1080x1920."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "structured-gaussian-splatting_b200"), ROOT]
from gsr_b200 import losses  # noqa: E402
from oracle import loss_oracle  # noqa: E402  (window taps only)


def torch_loss(x, t, win):
    C = x.shape[0]
    conv = lambda a: torch.nn.functional.conv2d(a[None], win, padding=5, groups=C)[0]
    mu1, mu2 = conv(x), conv(t)
    s1, s2, s12 = conv(x * x) - mu1 * mu1, conv(t * t) - mu2 * mu2, conv(x * t) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))
    return 0.8 * (x - t).abs().mean() + 0.2 * (1.0 - m.mean())


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    g = torch.Generator().manual_seed(0)
    x = torch.rand(3, 1080, 1920, generator=g).cuda().requires_grad_(True)
    t = torch.rand(3, 1080, 1920, generator=g).cuda()
    w1 = torch.from_numpy(loss_oracle.window_1d()).cuda()
    win = torch.outer(w1, w1)[None, None].expand(3, 1, 11, 11).contiguous()

    def fused():
        x.grad = None
        losses.l1_ssim_loss(x, t, 0.2).backward()

    def plain():
        x.grad = None
        torch_loss(x, t, win).backward()

    print(json.dumps({"fused_l1_ssim_ms": timeit(fused), "torch_l1_ssim_ms": timeit(plain),
                      "elements": x.numel(), "algorithmic_bytes": 44 * x.numel()}))


if __name__ == "__main__":
    main()
