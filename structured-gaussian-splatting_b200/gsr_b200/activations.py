"""The model's parameter activations in one launch per direction.

activate(_scaling, _rotation, _opacity) -> (scaling, rotation, opacity) is GaussianModel.get_scaling / get_rotation /
get_opacity (reference scene/gaussian_model.py:33-40,100-125: torch.exp, torch.nn.functional.normalize, torch.sigmoid)
with autograd support; csrc/activate.cu.  With diff_gaussian_rasterization.set_grad_mode("overwrite" | "accumulate") set,
leaf parameters with a preallocated .grad receive their gradients in place.  CUDA only; no CPU fallback.
"""
import torch

from . import load_binding


class _Activate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scaling, rotation, opacity):
        import diff_gaussian_rasterization as dgr
        s, q, o = load_binding().activate(scaling, rotation, opacity)
        ctx.save_for_backward(scaling, rotation, opacity)
        ctx.grad_mode = dgr._GRAD_MODE
        t = tuple(dgr._grad_target(x) for x in (scaling, rotation, opacity)) if ctx.grad_mode else None
        ctx.targets = t if (t is not None and all(x is not None for x in t)) else None
        return s, q, o

    @staticmethod
    def backward(ctx, g_s, g_q, g_o):
        scaling, rotation, opacity = ctx.saved_tensors
        z = lambda g, like, shape: torch.zeros(shape, device=like.device, dtype=torch.float32) if g is None else g
        P = scaling.shape[0]
        g_s, g_q, g_o = z(g_s, scaling, (P, 3)), z(g_q, scaling, (P, 4)), z(g_o, scaling, (P, 1))
        e = torch.empty(0)
        if ctx.targets is not None:
            load_binding().activate_backward(scaling, rotation, opacity, g_s, g_q, g_o, ctx.targets[0], ctx.targets[1], ctx.targets[2],
                                             ctx.grad_mode == "accumulate")
            return None, None, None
        ds, dq, do = load_binding().activate_backward(scaling, rotation, opacity, g_s, g_q, g_o, e, e, e, False)
        return ds.view_as(scaling), dq.view_as(rotation), do.view_as(opacity)


def activate(scaling: torch.Tensor, rotation: torch.Tensor, opacity: torch.Tensor):
    """(exp(scaling) [P,3], normalize(rotation) [P,4], sigmoid(opacity) [P,1])"""
    if not scaling.is_cuda:
        raise RuntimeError("gsr_b200.activations.activate: CUDA tensors only (no CPU fallback)")
    return _Activate.apply(scaling, rotation, opacity)
