"""CPU restatement (numpy, float64 accumulation) of the training loss around the rasterizer -- TEST INFRASTRUCTURE ONLY.

Follows the reference's utils/loss_utils.py:
  l1_loss  :17-18   mean |x - y|
  gaussian :23-25   11 taps exp(-(i-5)^2 / (2 1.5^2)), float32, normalised by their sum
  _ssim    :46-63   depthwise conv2d with the 11x11 outer-product window, zero padding 5, of x, y, x*x, y*y, x*y;
                    map = (2 mu1 mu2 + C1)(2 s12 + C2) / ((mu1^2 + mu2^2 + C1)(s1 + s2 + C2)), C1 = 0.01^2, C2 = 0.03^2; mean
and train.py:104-105  loss = (1 - lambda) l1 + lambda (1 - ssim).
The gradient is derived by hand (chain rule through the five windowed moments; the adjoint of the zero-padded symmetric
window is the window itself) and pinned against torch autograd over the reference functions by
tests/golden/loss_l1_ssim.npz (tests/golden/make_loss_golden.py).  Only tests/, smoke() and bench.py's CPU legs import it.
"""
from __future__ import annotations

import math

import numpy as np

WINDOW, SIGMA = 11, 1.5
C1, C2 = 0.01 ** 2, 0.03 ** 2


def window_1d() -> np.ndarray:
    g = np.array([math.exp(-(x - WINDOW // 2) ** 2 / float(2 * SIGMA ** 2)) for x in range(WINDOW)], np.float32)
    return (g / g.sum(dtype=np.float32)).astype(np.float32)


def _conv(img: np.ndarray) -> np.ndarray:
    """depthwise 11x11 window, zero padding 5; img [C,H,W] -> [C,H,W] (float64 accumulation)"""
    w1 = window_1d()
    w2 = np.outer(w1, w1).astype(np.float32).astype(np.float64)          # the reference's float 2-D window
    C, H, W = img.shape
    r = WINDOW // 2
    p = np.zeros((C, H + 2 * r, W + 2 * r), np.float64)
    p[:, r:r + H, r:r + W] = img
    out = np.zeros((C, H, W), np.float64)
    for dy in range(WINDOW):
        for dx in range(WINDOW):
            out += w2[dy, dx] * p[:, dy:dy + H, dx:dx + W]
    return out


def l1_ssim(image: np.ndarray, target: np.ndarray, lambda_dssim: float = 0.2, with_grad: bool = True):
    """returns dict(l1, ssim, loss[, grad]) for [C,H,W] float images"""
    x = image.astype(np.float64)
    y = target.astype(np.float64)
    n = x.size
    mu1, mu2 = _conv(x), _conv(y)
    xx, yy, xy = _conv(x * x), _conv(y * y), _conv(x * y)
    s1, s2, s12 = xx - mu1 * mu1, yy - mu2 * mu2, xy - mu1 * mu2
    a, b = 2 * mu1 * mu2 + C1, 2 * s12 + C2
    c, d = mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    m = a * b / (c * d)
    l1 = np.abs(x - y).mean()
    ssim = m.mean()
    out = {"l1": l1, "ssim": ssim, "loss": (1 - lambda_dssim) * l1 + lambda_dssim * (1 - ssim)}
    if with_grad:
        dmu1 = 2 * mu2 * (b - a) / (c * d) + 2 * mu1 * m * (1 / d - 1 / c)
        dxx = -m / d
        dxy = 2 * a / (c * d)
        dssim = _conv(dmu1) + 2 * x * _conv(dxx) + y * _conv(dxy)       # d(sum m)/dx
        out["grad"] = (1 - lambda_dssim) * np.sign(x - y) / n - lambda_dssim * dssim / n
    return out
