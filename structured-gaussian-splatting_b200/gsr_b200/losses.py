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
