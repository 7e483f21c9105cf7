"""Drop-in `diff_gaussian_rasterization` package backed by the B200-native core.

Public surface mirrors the reference package
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py):
  GaussianRasterizationSettings  (:157-169)  12-field NamedTuple
  GaussianRasterizer             (:171-220)  nn.Module: forward(...) -> (color[3,H,W], radii[P]), markVisible
  rasterize_gaussians            (:21-42)    functional form
  _RasterizeGaussians            (:44-155)   autograd.Function, same argument / gradient order
so gaussian_renderer/__init__.py, train.py, train_lgm.py and render.py call it unchanged.
"""
import os
import threading
import weakref
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C


def cpu_deep_copy_tuple(input_tuple):
    """Host-side snapshot of an argument tuple (used for the debug dumps)."""
    return tuple(x.cpu().clone() if isinstance(x, torch.Tensor) else x for x in input_tuple)


def _call_with_snapshot(fn, args, debug, dump_name, stage):
    """debug=True keeps the reference's behaviour: snapshot inputs first, dump them if the core raises."""
    if not debug:
        return fn(*args)
    saved = cpu_deep_copy_tuple(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_name)
        print("\nAn error occured in %s. Please forward %s for debugging." % (stage, dump_name))
        raise


# ---- extension (off by default): gradients straight into the leaf parameters' .grad buffers -------------------
# A multi-view step sums the parameter gradients of its views.  Through plain autograd every view costs one
# elementwise `grad += new` per parameter (3 x 236 B per Gaussian at SH degree 3) on top of the backward's own
# writes.  With a mode set here, an input that IS a leaf parameter with a preallocated contiguous fp32 .grad gets
# its gradient written (mode "overwrite") or added (mode "accumulate") by the backward kernels themselves, and
# autograd receives None for it.  Inputs that are not such leaves (the reference trainer passes activations)
# are unaffected, as is everything when the mode is None.
#
# Contract (checked in backward; a violated check falls back to handing the gradient to autograd, which is always right):
#   * the target is the leaf's .grad AS IT IS WHEN BACKWARD RUNS: if .grad was rebound or set to None between forward and
#     backward (optimizer.zero_grad(set_to_none=True)), nothing is written in place for that leaf;
#   * "overwrite" replaces what the buffer holds, so at most ONE render of a leaf may be pending in that mode: a second
#     forward of the same leaf before the first one's backward raises (it would silently drop the first view's gradient);
#     use "accumulate" for the other views of a step, as gsr_b200.dp.backward_views does;
#   * in "accumulate" mode dL/dmean3D is added with float atomics from two concurrently running kernels: sums are exact
#     to rounding but not bit-reproducible from run to run (like the reference's atomicAdd gradients).
_GRAD_MODE = None
_PENDING_OVERWRITE = {}          # data_ptr of a .grad buffer -> weak reference to the token of the pending "overwrite" render


_TLS = threading.local()


class _Token:
    """lives as long as the autograd node of a pending render (dropped graphs release their claim automatically)"""
    __slots__ = ("__weakref__",)


def set_grad_mode(mode):
    """None (reference behaviour), "overwrite", "accumulate" or "deferred"; returns the previous mode.
    "deferred": a step that renders several views of the same leaves runs, per view, only the blending backward; the
    per-Gaussian backward of all pending views (cov2D / cov3D / SH / projection: bandwidth-bound, 0.4 ms per view at C3) runs
    ONCE in flush_deferred(), reading the parameters once and writing every gradient row once (0.2 ms per view).  Until the
    flush the leaves' .grad do not contain the pending views; means2D.grad of every view is set by the flush."""
    global _GRAD_MODE
    if mode not in (None, "overwrite", "accumulate", "deferred"):
        raise ValueError("grad mode must be None, 'overwrite', 'accumulate' or 'deferred'")
    prev, _GRAD_MODE = _GRAD_MODE, mode
    return prev


_DEFERRED = []                   # pending views of the "deferred" mode (one record per backward)


def pending_deferred():
    return len(_DEFERRED)


def flush_deferred(accumulate=False, n_chunks=0, after_chunk=None):
    """Per-Gaussian backward of all pending "deferred" views, summed into the leaves' .grad buffers (replacing their contents
    unless `accumulate`).  n_chunks > 0: in ranges of Gaussians, `after_chunk(g_begin, g_end)` being called on the host after
    each range has been queued (gsr_b200.dp starts that range's all-reduce from it).  Returns the number of views flushed."""
    n = len(_DEFERRED)
    if n == 0:
        return 0
    recs, first = list(_DEFERRED), _DEFERRED[0]
    del _DEFERRED[:]
    for r in recs:
        if any(a is not b for a, b in zip(r["leaves"], first["leaves"])) or r["key"] != first["key"]:
            raise RuntimeError("flush_deferred: the pending views do not share parameters / raster settings")
    means3D, opacities, sh, scales, rotations = first["leaves"]
    targets = []
    for leaf in first["leaves"]:
        g = _grad_target(leaf)
        if g is None:
            raise RuntimeError("flush_deferred: a parameter lost its preallocated .grad buffer since the views were rendered")
        targets.append(g)
    H, W, tanfovx, tanfovy, degree, scale_modifier = first["key"]
    for base in range(0, n, 16):
        part = recs[base:base + 16]
        with torch.no_grad():
            m2d = _C.rasterize_gaussians_backward_views(
                means3D, scales, rotations, scale_modifier, sh, degree, tanfovx, tanfovy, H, W,
                [r["viewmatrix"] for r in part], [r["projmatrix"] for r in part], [r["campos"] for r in part],
                [r["radii"] for r in part], [r["geom"] for r in part], [r["accum"] for r in part],
                targets[0], targets[1], targets[2], targets[3], targets[4], bool(accumulate or base > 0), int(n_chunks), after_chunk)
        for r, g2 in zip(part, m2d):
            m = r["means2D"]
            if m is not None and m.requires_grad and m.is_leaf:
                m.grad = g2 if m.grad is None else m.grad + g2
    return n


def _grad_target(t):
    g = getattr(t, "grad", None)
    if (t.numel() != 0 and t.is_leaf and t.requires_grad and g is not None and g.is_cuda and g.dtype == torch.float32
            and g.is_contiguous() and g.numel() == t.numel()):
        return g
    return None


# ---- extension (off by default): the per-Gaussian backward of a view in ranges of Gaussians, with a host callback after
# each range (gsr_backward_staged): gsr_b200.dp starts the all-reduce of a range's gradients from it.  Only honoured
# together with a grad mode (the gradients must land in caller-owned buffers for the callback to see them).
_CHUNK_HOOK = None               # (n_chunks, callable(g_begin, g_end)) or None


def set_chunk_hook(hook):
    """hook = (n_chunks, callback) or None; returns the previous hook."""
    global _CHUNK_HOOK
    prev, _CHUNK_HOOK = _CHUNK_HOOK, hook
    return prev


# ---- extension (off by default): the front of a view (preprocess, depth order, tile split) on a stream of its own
# (gsr_forward_front).  Those stages are bound by memory latency and bandwidth, the blending kernels by instruction issue: a
# step that renders several views issues the forward of view i+1 BEFORE the backward of view i, and with a high-priority front
# stream the next view's front runs inside the current view's blending kernels (gsr_b200.dp.deferred_views(pipeline=True)).
# Contract: the front stream does NOT wait for the stream the render is called on - that is the point - so every input of
# the render (parameters, camera tensors) must be valid on it.  set_front_stream() makes it wait for the calling stream ONCE,
# at the moment of the call (so: call it after the optimizer step, before the first render of the step); inputs produced
# later on another stream (a camera copied in by a loader) are the caller's to order: front.wait_event(...).
_FRONT = None                    # torch.cuda.Stream or None
_FRONT_ZERO_ACCUM = os.environ.get("GSR_FRONT_ZERO_ACCUM", "1") != "0"


def set_front_stream(stream, wait=True):
    """stream: a torch.cuda.Stream (ideally created with a higher priority than the rendering stream) or None; returns the
    previous one.  The stream waits for everything queued so far on the current stream unless wait=False (a caller that
    alternates between front streams inside a step orders each of them once, at the start of the step)."""
    global _FRONT
    prev, _FRONT = _FRONT, stream
    if stream is not None and wait:
        stream.wait_stream(torch.cuda.current_stream(stream.device))
    return prev


def get_front_stream():
    """The stream set by set_front_stream (None when none is set)."""
    return _FRONT


def front_stream(device=None, high_priority=True, index=0):
    """A stream for set_front_stream (one per device, priority class and index, created on first use); high priority by
    default, so the front's small kernels get SM slots as the blending kernels' thread blocks retire."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    key = (dev, bool(high_priority), int(index))
    st = _FRONT_STREAMS.get(key)
    if st is None:
        prio = 0
        if high_priority:
            lo, hi = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
            prio = min(lo, hi)
        st = _FRONT_STREAMS[key] = torch.cuda.Stream(device=dev, priority=prio)
    return st


_FRONT_STREAMS = {}


# ---- extension (off by default): the preprocess stage of a step's views in ONE pass over the parameters
# (gsr_preprocess_views).  A view's preprocess moves ~300 bytes per Gaussian, 236 of them the parameters (192 the SH row);
# a step that renders V views of the same Gaussians reads them V times.  prepare_views() runs that stage for up to 16
# cameras at once - each thread keeps its Gaussian and walks the cameras - and parks every view's state; the following
# GaussianRasterizer calls with those cameras find it and run only depth order, tile split and blending.  Results are those
# of the plain call (the same device function computes them).
# A parked view is matched by the IDENTITY and VERSION of the tensors it was prepared from (parameters and camera): a
# parameter updated in place, a camera buffer rewritten, another raster setting - and the render takes the plain path.
_PREPARED = {}


def _tensor_key(t):
    return (t.data_ptr(), t._version, tuple(t.shape))


def _prepared_key(means3D, opacities, sh, scales, rotations, rs):
    return (_tensor_key(means3D), _tensor_key(opacities), _tensor_key(sh), _tensor_key(scales), _tensor_key(rotations),
            _tensor_key(rs.viewmatrix), _tensor_key(rs.projmatrix), _tensor_key(rs.campos), int(rs.image_height), int(rs.image_width),
            float(rs.tanfovx), float(rs.tanfovy), int(rs.sh_degree), float(rs.scale_modifier), bool(rs.prefiltered))


def prepare_views(means3D, opacities, shs, scales, rotations, settings):
    """Preprocesses the views of `settings` (a list of GaussianRasterizationSettings sharing image size, field of view, SH
    degree and scale modifier) in one pass over the parameters; the renders that follow with exactly these tensors and
    settings use the result.  Runs on the front stream when one is set (set_front_stream), else on the current stream.
    Views prepared earlier and never rendered are dropped.  Returns the number of views prepared."""
    _PREPARED.clear()
    settings = list(settings)
    if not settings or not means3D.is_cuda or means3D.numel() == 0 or shs is None or shs.dim() != 3:
        return 0
    first = settings[0]
    same = lambda r: (r.image_height, r.image_width, r.tanfovx, r.tanfovy, r.sh_degree, r.scale_modifier, r.prefiltered)
    if any(same(r) != same(first) for r in settings) or any(r.debug for r in settings):
        return 0
    tensors = [means3D, opacities, shs, scales, rotations] + [t for r in settings for t in (r.viewmatrix, r.projmatrix, r.campos)]
    if not all(t.is_cuda and t.is_contiguous() and t.dtype == torch.float32 for t in tensors):
        return 0
    front = _FRONT if (_FRONT is not None and _FRONT.device == means3D.device) else None
    stream = front if front is not None else torch.cuda.current_stream(means3D.device)
    n = 0
    with torch.no_grad(), torch.cuda.stream(stream):
        for base in range(0, len(settings), 16):
            part = settings[base:base + 16]
            geoms, radii = _C.prepare_views(means3D, opacities, scales, rotations, first.scale_modifier, shs, first.sh_degree,
                                            first.tanfovx, first.tanfovy, first.image_height, first.image_width,
                                            [r.viewmatrix for r in part], [r.projmatrix for r in part], [r.campos for r in part],
                                            bool(first.prefiltered))
            ev = torch.cuda.Event()
            ev.record(stream)
            for r, g, rad in zip(part, geoms, radii):
                _PREPARED[_prepared_key(means3D, opacities, shs, scales, rotations, r)] = (g, rad, ev)
                n += 1
    return n


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings):
        rs = raster_settings
        # argument order of _C.rasterize_gaussians (rasterize_points.h:18-38)
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh, rs.sh_degree,
                rs.campos, rs.prefiltered, rs.debug)
        front = _FRONT
        use_front = front is not None and means3D.is_cuda and front.device == means3D.device and not rs.debug
        # deferred steps with a front stream: the accumulator rows the blending backward adds into are cleared on the front
        # stream, ahead of the front's kernels, instead of in front of the blending backward (144 MB at C3)
        ctx.zeroed_accum = None
        if (_FRONT_ZERO_ACCUM and use_front and _GRAD_MODE == "deferred" and getattr(_TLS, "grad_enabled", False)
                and means3D.shape[0] > 0):
            with torch.cuda.stream(front):
                ctx.zeroed_accum = torch.zeros((means3D.shape[0], 12), dtype=torch.float32, device=means3D.device)
            ctx.zeroed_accum.record_stream(torch.cuda.current_stream(means3D.device))
        parked = None
        if _PREPARED and means3D.is_cuda and colors_precomp.numel() == 0 and cov3Ds_precomp.numel() == 0 and sh.dim() == 3:
            parked = _PREPARED.pop(_prepared_key(means3D, opacities, sh, scales, rotations, rs), None)
        if parked is not None:
            geom, rad, ev = parked
            (front if use_front else torch.cuda.current_stream(means3D.device)).wait_event(ev)
            num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _C.rasterize_gaussians_prepared(
                *args, geom, rad, front.cuda_stream if use_front else 0)
        elif use_front:
            num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _C.rasterize_gaussians_front(*args, front.cuda_stream)
        else:
            num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _call_with_snapshot(
                _C.rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.grad_mode = _GRAD_MODE
        if _GRAD_MODE is not None:
            ctx.grad_leaves = (means3D, opacities, sh, scales, rotations)
            ctx.grad_targets = tuple(_grad_target(t) for t in ctx.grad_leaves)
            ctx.means2D_ref = means2D if _GRAD_MODE == "deferred" else None
            # (autograd runs forward() with grad mode off: the caller's mode was noted by rasterize_gaussians())
            if _GRAD_MODE == "overwrite" and getattr(_TLS, "grad_enabled", False):
                ctx.overwrite_token = _Token()
                for g in ctx.grad_targets:
                    if g is not None:
                        other = _PENDING_OVERWRITE.get(g.data_ptr())
                        if other is not None and other() is not None:
                            raise RuntimeError("set_grad_mode('overwrite'): this parameter already has a pending render whose "
                                               "backward will overwrite its .grad; use 'accumulate' for further views")
                        _PENDING_OVERWRITE[g.data_ptr()] = weakref.ref(ctx.overwrite_token)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        # no zero-filled stand-in for the gradient of `radii` (an int32 [P] fill per backward call: 12 MB at C3)
        ctx.set_materialize_grads(False)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii):
        rs = ctx.raster_settings
        if grad_out_color is None:              # (only with set_materialize_grads(False): the image took no part in the loss)
            grad_out_color = torch.zeros((3, rs.image_height, rs.image_width), dtype=torch.float32, device=ctx.saved_tensors[1].device)
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer, imgBuffer = \
            ctx.saved_tensors
        # argument order of _C.rasterize_gaussians_backward (rasterize_points.h:40-62)
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix,
                rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, sh, rs.sh_degree, rs.campos, geomBuffer,
                ctx.num_rendered, binningBuffer, imgBuffer, rs.debug)
        deferred = (ctx.grad_mode == "deferred" and colors_precomp.numel() == 0 and cov3Ds_precomp.numel() == 0 and sh.dim() == 3
                    and all(g is not None and getattr(leaf, "grad", None) is g for g, leaf in zip(ctx.grad_targets, ctx.grad_leaves)))
        if deferred:
            # the blending backward now, the per-Gaussian backward with the other pending views in flush_deferred()
            accum = _C.rasterize_gaussians_backward_blend(rs.bg, means3D, radii, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                                                          grad_out_color, int(sh.shape[1]), rs.sh_degree, geomBuffer, ctx.num_rendered,
                                                          binningBuffer, imgBuffer, rs.debug,
                                                          ctx.zeroed_accum if ctx.zeroed_accum is not None else torch.Tensor([]))
            ctx.zeroed_accum = None
            _DEFERRED.append(dict(accum=accum, geom=geomBuffer, radii=radii, viewmatrix=rs.viewmatrix, projmatrix=rs.projmatrix,
                                  campos=rs.campos, means2D=ctx.means2D_ref, leaves=ctx.grad_leaves,
                                  key=(rs.image_height, rs.image_width, rs.tanfovx, rs.tanfovy, rs.sh_degree, rs.scale_modifier)))
            return (None,) * 9
        if ctx.grad_mode is None or ctx.grad_mode == "deferred":
            (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
             grad_rotations) = _call_with_snapshot(_C.rasterize_gaussians_backward, args, rs.debug, "snapshot_bw.dump", "backward")
        else:
            empty = torch.Tensor([])
            if ctx.grad_mode == "overwrite":
                for g in ctx.grad_targets:
                    if g is not None:
                        _PENDING_OVERWRITE.pop(g.data_ptr(), None)
            # the target must still be the leaf's .grad: a buffer rebound or dropped since the forward gets nothing in place
            targets = tuple(g if (g is not None and getattr(leaf, "grad", None) is g) else None
                            for g, leaf in zip(ctx.grad_targets, ctx.grad_leaves))
            into = tuple(empty if t is None else t for t in targets)
            hook = _CHUNK_HOOK
            if hook is not None and all(t is not None for t in targets):
                (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
                 grad_rotations) = _call_with_snapshot(_C.rasterize_gaussians_backward_chunked,
                                                       args + into + (ctx.grad_mode == "accumulate", int(hook[0]), hook[1]),
                                                       rs.debug, "snapshot_bw.dump", "backward")
            else:
                (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
                 grad_rotations) = _call_with_snapshot(_C.rasterize_gaussians_backward_into,
                                                       args + into + (ctx.grad_mode == "accumulate",), rs.debug,
                                                       "snapshot_bw.dump", "backward")
            t_m3, t_op, t_sh, t_sc, t_rt = targets
            # written in place by the kernels: autograd gets None for these inputs
            grad_means3D = None if t_m3 is not None else grad_means3D
            grad_opacities = None if t_op is not None else grad_opacities
            grad_sh = None if t_sh is not None else grad_sh
            grad_scales = None if t_sc is not None else grad_scales
            grad_rotations = None if t_rt is not None else grad_rotations

        def present(g, inp):      # absent inputs (0-element sentinels) get no gradient
            return g if (g is not None and inp.numel() != 0) else None

        return (grad_means3D, grad_means2D, present(grad_sh, sh), present(grad_colors_precomp, colors_precomp),
                grad_opacities, present(grad_scales, scales), present(grad_rotations, rotations),
                present(grad_cov3Ds_precomp, cov3Ds_precomp), None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    _TLS.grad_enabled = torch.is_grad_enabled()
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                                     raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """bool[P]: view-space z > 0.2 for the settings' camera."""
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        has_sr_any = scales is not None or rotations is not None
        has_sr_all = scales is not None and rotations is not None
        if (not has_sr_all and cov3D_precomp is None) or (has_sr_any and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        absent = torch.Tensor([])      # the reference's "not given" sentinel: an empty CPU tensor
        return rasterize_gaussians(
            means3D, means2D,
            absent if shs is None else shs,
            absent if colors_precomp is None else colors_precomp,
            opacities,
            absent if scales is None else scales,
            absent if rotations is None else rotations,
            absent if cov3D_precomp is None else cov3D_precomp,
            self.raster_settings)
