"""Fused image losses of the training step around the rasterizer.

l1_loss(network_output, gt) is the drop-in for the reference's utils/loss_utils.py:17-18
(`torch.abs(network_output - gt).mean()`): same value, same gradient (sign(x - gt) / n, 0 where equal), but the
forward and the gradient come out of ONE kernel that reads both images once (csrc/loss.cu) instead of six
elementwise launches and three full-size temporaries.  CUDA tensors only; there is no CPU fallback.
"""
import torch

from . import load_binding


class _L1Loss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, target):
        s, grad = load_binding().l1_loss(image, target)
        ctx.save_for_backward(grad)
        return (s / image.numel()).reshape(())

    @staticmethod
    def backward(ctx, grad_out):
        (grad,) = ctx.saved_tensors
        return grad * grad_out, None


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    if not network_output.is_cuda:
        raise RuntimeError("gsr_b200.losses.l1_loss: CUDA tensors only (no CPU fallback)")
    return _L1Loss.apply(network_output, gt)


class _L1SsimLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, target, lambda_dssim):
        need_grad = image.requires_grad
        sums, grad = load_binding().l1_ssim_loss(image, target, float(lambda_dssim), need_grad)
        ctx.save_for_backward(grad)
        n = image.numel()
        return ((1.0 - lambda_dssim) * sums[0] / n + lambda_dssim * (1.0 - sums[1] / n)).reshape(())

    @staticmethod
    def backward(ctx, grad_out):
        (grad,) = ctx.saved_tensors
        return grad * grad_out, None, None


def l1_ssim_loss(image: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2) -> torch.Tensor:
    """(1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt)): the training loss of train.py:104-105
    (lambda_dssim = 0.2, arguments/__init__.py), value and dL/dimage from two kernels (csrc/ssim.cu).  [C,H,W] images."""
    if not image.is_cuda:
        raise RuntimeError("gsr_b200.losses.l1_ssim_loss: CUDA tensors only (no CPU fallback)")
    return _L1SsimLoss.apply(image, gt, lambda_dssim)


def ssim(img1: torch.Tensor, img2: torch.Tensor) -> torch.Tensor:
    """Mean SSIM of two [C,H,W] images (utils/loss_utils.py:38-63, window 11, size_average=True), differentiable in img1."""
    if not img1.is_cuda:
        raise RuntimeError("gsr_b200.losses.ssim: CUDA tensors only (no CPU fallback)")
    return 1.0 - _L1SsimLoss.apply(img1, img2, 1.0)
