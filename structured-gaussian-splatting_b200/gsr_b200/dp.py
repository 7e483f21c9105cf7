"""View-sharded data-parallel rasterisation step (BASELINE.json config 5).

One process per GPU (torchrun).  Every rank holds the full Gaussian replica, renders its slice of the
step's cameras (view k goes to rank k mod world), accumulates parameter gradients locally in ONE flat
fp32 buffer, and the buffer is summed across ranks with a single NCCL all-reduce over NVLink/NVSwitch.
The reference has no multi-GPU code; this is the layer its train.py would call instead of a
single-view `render() ... loss.backward()`.

Flat layout (floats per Gaussian: 3 + 3 + 4 + 1 + 3*M):
    means3D P*3 | scales P*3 | rotations P*4 | opacities P*1 | shs P*M*3
Each parameter's .grad is a view into the buffer, so autograd accumulates in place and no
flatten / unflatten copy happens before the collective.

The collective ("bucketed ... issued from the K10 epilogue of the last local view on a side stream"):
no gradient is final before the rank's last view has gone through its per-Gaussian backward, so that stage of the
LAST view runs in ranges of Gaussians (gsr_backward_staged) and the all-reduce of a range - one coalesced NCCL call
over the range's rows of every parameter - is issued as soon as the range has been queued; NCCL's stream waits for
exactly that point of the compute stream, so range k travels over NVLink while range k+1 is computed.  The buffer can
be allocated from NCCL's own allocator and registered with the communicator (ncclMemAlloc / ncclCommRegister through
torch.cuda.MemPool), which lets NCCL run its NVLS (in-switch reduction) path without staging copies.
"""
from __future__ import annotations

from typing import Callable, Dict, Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist

PARAM_ORDER = ("means3D", "scales", "rotations", "opacities", "shs")


def init_nccl(device, high_priority: bool = True, **kw):
    """dist.init_process_group("nccl") with NCCL's kernels on a HIGH-PRIORITY stream.  The overlapped steps below issue the
    all-reduce of a range of Gaussians while the next range of the per-Gaussian backward fills every SM; an NCCL thread
    block needs most of an SM to itself, so at default priority it only gets one in the compute kernels' tail (measured at
    C3, N = 2: 17.85 -> 17.32 ms per step).  Extra keyword arguments go to init_process_group."""
    opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True) if high_priority else None
    return dist.init_process_group("nccl", device_id=device, pg_options=opts, **kw)


def shard_views(n_views: int, rank: int, world: int) -> List[int]:
    """Views owned by `rank`: k with k mod world == rank."""
    return [k for k in range(n_views) if k % world == rank]


class FlatGrads:
    """Owns the flat gradient buffer and binds each parameter's .grad to its slice."""

    def __init__(self, params: Dict[str, torch.Tensor], nccl_registered: bool = False, group=None):
        self.params = params
        self.names = [n for n in PARAM_ORDER if n in params]
        total = sum(params[n].numel() for n in self.names)
        ref = params[self.names[0]]
        self.P = int(ref.shape[0])
        self.registered = False
        self.flat = None
        if nccl_registered and ref.is_cuda and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            # buffer from NCCL's allocator, registered with the communicator: zero-copy NVLS all-reduce
            try:
                pg = group if group is not None else dist.distributed_c10d._get_default_group()
                backend = pg._get_backend(ref.device)
                self._backend = backend
                self._pool = torch.cuda.MemPool(backend.mem_allocator)
                with torch.cuda.use_mem_pool(self._pool):
                    self.flat = torch.zeros(total, dtype=torch.float32, device=ref.device)
                backend.register_mem_pool(self._pool)
                self.registered = True
            except Exception as ex:      # older NCCL / no NVLS: a plain buffer works everywhere
                self.flat = None
                self.register_error = str(ex)[:200]
        if self.flat is None:
            self.flat = torch.zeros(total, dtype=torch.float32, device=ref.device)
        self.slices = {}
        o = 0
        for n in self.names:
            k = params[n].numel()
            self.slices[n] = (o, o + k)
            params[n].grad = self.flat[o:o + k].view_as(params[n])
            o += k

    def zero_(self):
        self.flat.zero_()

    def release(self):
        """Drops the buffer (the parameters' .grad become None).  A buffer registered with NCCL MUST be released before the
        process group is destroyed: its memory pool deregisters from the communicator when it is freed."""
        for n in self.names:
            self.params[n].grad = None
        self.flat = None
        if self.registered:
            try:
                self._backend.deregister_mem_pool(self._pool)
            except Exception:
                pass
            self._pool = None
            self._backend = None
            self.registered = False

    def grad(self, name: str) -> torch.Tensor:
        a, b = self.slices[name]
        return self.flat[a:b].view_as(self.params[name])

    def nbytes(self) -> int:
        return self.flat.numel() * 4

    def all_reduce(self, group=None, async_op: bool = False):
        """Sum over ranks (no-op for world size 1)."""
        if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
            return None
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)

    def row_slices(self, g0: int, g1: int) -> List[torch.Tensor]:
        """The rows [g0, g1) of every parameter's gradient, as views into the flat buffer."""
        out = []
        for n in self.names:
            a, b = self.slices[n]
            k = (b - a) // self.P
            out.append(self.flat[a + g0 * k:a + g1 * k])
        return out

    def all_reduce_rows(self, g0: int, g1: int, group=None):
        """One coalesced all-reduce (a single NCCL group call) over the rows [g0, g1) of every parameter; asynchronous:
        NCCL's stream waits for the current stream as of this call.  Returns an object with .wait(), or None."""
        if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
            return None
        ts = self.row_slices(g0, g1)
        with dist._coalescing_manager(group=group, device=self.flat.device, async_ops=True) as cm:
            for t in ts:
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return cm


def make_leaf_params(scene: Dict[str, torch.Tensor], device) -> Dict[str, torch.Tensor]:
    return {k: scene[k].to(device).clone().requires_grad_(True) for k in PARAM_ORDER if k in scene}


def view_settings(rasterizer_mod, cam, bg, sh_degree: int, scale_modifier: float = 1.0):
    """GaussianRasterizationSettings of one camera, as gaussian_renderer.render() builds them (gaussian_renderer/__init__.py:36-50)."""
    return rasterizer_mod.GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=bg,
        scale_modifier=scale_modifier, viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix, sh_degree=sh_degree,
        campos=cam.campos, prefiltered=False, debug=False)


def render_view(rasterizer_mod, params, cam, bg, sh_degree: int, scale_modifier: float = 1.0):
    """One forward through the drop-in API, exactly as gaussian_renderer.render() calls it
    (gaussian_renderer/__init__.py:36-93).  Returns (color[3,H,W], radii, means2D grad sink)."""
    rs = view_settings(rasterizer_mod, cam, bg, sh_degree, scale_modifier)
    means2D = torch.zeros_like(params["means3D"], requires_grad=True)
    color, radii = rasterizer_mod.GaussianRasterizer(rs)(
        means3D=params["means3D"], means2D=means2D, opacities=params["opacities"], shs=params["shs"],
        scales=params["scales"], rotations=params["rotations"])
    return color, radii, means2D


def prepare_step_views(rasterizer_mod, params, cams: Sequence, bg, sh_degree: int, view_ids: Iterable[int]) -> int:
    """The preprocess stage of the step's views in one pass over the parameters (diff_gaussian_rasterization.prepare_views /
    gsr_preprocess_views): the render_view calls that follow with these cameras skip it.  Returns the number of views prepared
    (0 where the rasterizer module has no such entry point, e.g. the reference package)."""
    if not hasattr(rasterizer_mod, "prepare_views"):
        return 0
    return rasterizer_mod.prepare_views(params["means3D"], params["opacities"], params["shs"], params["scales"], params["rotations"],
                                        [view_settings(rasterizer_mod, cams[k], bg, sh_degree) for k in view_ids])


_STREAMS: List["torch.cuda.Stream"] = []


def _side_streams(n: int):
    while len(_STREAMS) < n:
        _STREAMS.append(torch.cuda.Stream())
    return _STREAMS[:n]


def backward_views(rasterizer_mod, params, flat: FlatGrads, cams: Sequence, bg, sh_degree: int,
                   loss_fn: Callable[[torch.Tensor, int], torch.Tensor], view_ids: Iterable[int], fused: bool = True,
                   n_streams: int = 1):
    """Renders the given views and back-propagates `loss_fn(color, view_id)` of each into the flat buffer.
    fused=True: the backward kernels write the first view's gradients straight into the buffer and ADD the following
    views' (diff_gaussian_rasterization.set_grad_mode) - no zero-fill, no per-view elementwise adds.
    fused=False: plain autograd accumulation into the zeroed buffer (what the reference API alone offers).
    n_streams > 1 (fused only): consecutive views alternate over that many CUDA streams, so the latency-bound sort stages
    and the HBM-bound per-Gaussian kernels of one view run next to the issue-bound blending kernels of another; the
    gradient accumulation stays ordered (the core fences its read-modify-write kernels in call order)."""
    total = None
    views = list(view_ids)
    if not views or not fused or not hasattr(rasterizer_mod, "set_grad_mode"):
        flat.zero_()
        fused = False
    if fused and n_streams > 1 and len(views) > 1 and params[flat.names[0]].is_cuda:
        cur = torch.cuda.current_stream()
        streams = _side_streams(n_streams)
        for st in streams:
            st.wait_stream(cur)
        losses = []
        for i, k in enumerate(views):
            with torch.cuda.stream(streams[i % n_streams]):
                prev = rasterizer_mod.set_grad_mode("overwrite" if i == 0 else "accumulate")
                try:
                    color, _radii, _m2d = render_view(rasterizer_mod, params, cams[k], bg, sh_degree)
                    loss = loss_fn(color, k)
                    loss.backward()
                finally:
                    rasterizer_mod.set_grad_mode(prev)
                losses.append(loss.detach())
        for st in streams:
            cur.wait_stream(st)
        for l in losses:
            l.record_stream(cur)
        return torch.stack(losses).sum()
    for i, k in enumerate(views):
        prev = rasterizer_mod.set_grad_mode(("overwrite" if i == 0 else "accumulate")) if fused else None
        try:
            color, _radii, _m2d = render_view(rasterizer_mod, params, cams[k], bg, sh_degree)
            loss = loss_fn(color, k)
            loss.backward()
        finally:
            if fused:
                rasterizer_mod.set_grad_mode(prev)
        total = loss.detach() if total is None else total + loss.detach()
    return total


def deferred_views(rasterizer_mod, params, flat: FlatGrads, cams: Sequence, bg, sh_degree: int,
                   loss_fn: Callable[[torch.Tensor, int], torch.Tensor], view_ids: Iterable[int], group=None,
                   overlap_chunks: int = 0, all_reduce: bool = True, pipeline: bool = False, prepare: bool = False,
                   front_streams: int = 2):
    """The step in DEFERRED form (diff_gaussian_rasterization.set_grad_mode("deferred")): every view runs its forward and its
    blending backward; the per-Gaussian backward of all the rank's views then runs once (parameters read once, every gradient
    row written once: 0.2 instead of 0.4 ms per view at C3), in `overlap_chunks` ranges of Gaussians when > 0, the all-reduce
    of a range being issued right behind it - the whole per-Gaussian stage of the step (not just the last view's) is the
    window the collective hides in.  Returns the summed local loss.
    pipeline=True: the forward of view i+1 is issued BEFORE the backward of view i, and the front of every forward
    (preprocess, depth order, tile split: latency- and bandwidth-bound) goes to a high-priority stream of its own
    (set_front_stream / gsr_forward_front), so it runs inside the issue-bound blending kernels of the views around it.
    Same results; two views' state is alive at a time.
    front_streams = n > 1 (with pipeline): the views' fronts alternate over n high-priority streams, so the (latency-bound)
    fronts of consecutive views may run next to each other: 1.962 against 1.975 ms per view at C3 for n = 2.
    prepare=True: the preprocess stage of all the step's views runs first, in one pass over the parameters
    (prepare_step_views): 236 of the ~300 bytes per Gaussian a view's preprocess moves are read once per step."""
    total = None
    views = list(view_ids)
    prev = rasterizer_mod.set_grad_mode("deferred")
    piped = (pipeline and len(views) > 1 and hasattr(rasterizer_mod, "set_front_stream")
             and params[flat.names[0]].is_cuda)
    fronts = []
    prev_front = None
    if piped:
        dev = params[flat.names[0]].device
        fronts = [rasterizer_mod.front_stream(dev, True, j) for j in range(max(1, front_streams))]
        # every front stream is ordered behind what is queued so far (the parameters); inside the step the views alternate
        # between them without further waits, so the fronts of consecutive views may run next to each other
        prev_front = rasterizer_mod.set_front_stream(fronts[-1])
        for f in reversed(fronts[:-1]):
            rasterizer_mod.set_front_stream(f)
    use_front = (lambda i: rasterizer_mod.set_front_stream(fronts[i % len(fronts)], wait=False)) if len(fronts) > 1 else (lambda i: None)
    try:
        if prepare and views:
            prepare_step_views(rasterizer_mod, params, cams, bg, sh_degree, views)
            if len(fronts) > 1:
                for f in fronts[1:]:
                    f.wait_stream(fronts[0])       # (the pass ran on the first front stream)
        if piped:
            use_front(0)
            ahead = render_view(rasterizer_mod, params, cams[views[0]], bg, sh_degree)
            for i, k in enumerate(views):
                color = ahead[0]
                use_front(i + 1)
                ahead = render_view(rasterizer_mod, params, cams[views[i + 1]], bg, sh_degree) if i + 1 < len(views) else None
                loss = loss_fn(color, k)
                loss.backward()
                total = loss.detach() if total is None else total + loss.detach()
        else:
            for k in views:
                color, _radii, _m2d = render_view(rasterizer_mod, params, cams[k], bg, sh_degree)
                loss = loss_fn(color, k)
                loss.backward()
                total = loss.detach() if total is None else total + loss.detach()
    finally:
        rasterizer_mod.set_grad_mode(prev)
        if piped:
            rasterizer_mod.set_front_stream(prev_front)
    multi = all_reduce and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    works = []
    if multi and overlap_chunks > 0:
        rasterizer_mod.flush_deferred(False, overlap_chunks, lambda g0, g1: works.append(flat.all_reduce_rows(g0, g1, group)))
        for w in works:
            if w is not None:
                w.wait()
    else:
        rasterizer_mod.flush_deferred(False)
        if multi:
            flat.all_reduce(group)
    return total


def dp_step(rasterizer_mod, params, flat: FlatGrads, cams: Sequence, bg, sh_degree: int,
            loss_fn: Callable[[torch.Tensor, int], torch.Tensor], view_ids: Iterable[int], group=None, fused: bool = True,
            n_streams: int = 1, overlap_chunks: int = 0):
    """Renders this rank's views, back-propagates `loss_fn(color, view_id)` for each into the flat
    buffer, then all-reduces it.  Returns the summed local loss (a device scalar).
    overlap_chunks = K > 0 (fused, one stream): the per-Gaussian backward of the LAST view runs in K ranges of Gaussians and
    each range's all-reduce is issued right behind it (see the module docstring); the result is the same sum."""
    views = list(view_ids)
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    if (overlap_chunks <= 0 or not multi or not fused or n_streams > 1 or not views
            or not hasattr(rasterizer_mod, "set_chunk_hook")):
        total = backward_views(rasterizer_mod, params, flat, cams, bg, sh_degree, loss_fn, views, fused, n_streams)
        flat.all_reduce(group)
        return total
    total = None
    if len(views) > 1:
        total = backward_views(rasterizer_mod, params, flat, cams, bg, sh_degree, loss_fn, views[:-1], True, 1)
    works = []
    prev_hook = rasterizer_mod.set_chunk_hook((overlap_chunks, lambda g0, g1: works.append(flat.all_reduce_rows(g0, g1, group))))
    prev_mode = rasterizer_mod.set_grad_mode("accumulate" if len(views) > 1 else "overwrite")
    try:
        k = views[-1]
        color, _radii, _m2d = render_view(rasterizer_mod, params, cams[k], bg, sh_degree)
        loss = loss_fn(color, k)
        loss.backward()
    finally:
        rasterizer_mod.set_grad_mode(prev_mode)
        rasterizer_mod.set_chunk_hook(prev_hook)
    if not works:                       # the hook was not honoured (a parameter without an in-place gradient target)
        flat.all_reduce(group)
    for w in works:
        if w is not None:
            w.wait()
    total = loss.detach() if total is None else total + loss.detach()
    return total
