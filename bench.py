#!/usr/bin/env python
"""bench.py - headline benchmark of the B200-native Gaussian rasterizer hot path.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json metric: "fwd+bwd rasterize ms/view @1080p 3M Gaussians; train views/s at 1-8 GPUs"):
config C3 - 3,000,000 synthetic Gaussians, SH degree 3, 1920x1080, cameras on the SURVEY 8(d) ring.
A step = every rank renders `--views-per-step` views forward+backward through the drop-in
GaussianRasterizer API over the full (replicated) Gaussian set, gradients accumulating in one flat
buffer, followed for N>1 by the NCCL all-reduce of that buffer (weak scaling: per-GPU work is fixed;
the last view's per-Gaussian backward runs in ranges and each range's all-reduce starts behind it).
value = total views/s over all ranks (ms/view on one GPU = 1000 / value at N=1).
The same step on config C5 (6 M Gaussians, 8 views per GPU and step = 64 views per step on 8 GPUs) is measured at
every N and reported in the `c5` object, so the C5 scaling curve can be read off the same lines.

Prints ONE JSON line (rank 0).  `--impl reference` times the UNMODIFIED reference CUDA extension (oracle/_ref, built
from /root/reference by oracle/build_ref.sh) on the same GPU, same scene, cameras and views per step, through plain
autograd - the optimisation bar SURVEY 8(d) names; the reference has no CPU implementation, so where that extension
or a GPU is missing (and for the `cpu_baseline` object) the reference algorithm's CPU restatement (oracle/, all host
threads) is timed on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "structured-gaussian-splatting_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "fwd+bwd rasterize views/s @1080p 3M Gaussians SH3 (ms/view = 1000*n_gpus/value)"
UNIT = "views/s"
WORKLOAD = "C3: 3M Gaussians, SH degree 3, 1920x1080, fwd+bwd via GaussianRasterizer API"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--views-per-step", type=int, default=8,
                    help="views each rank renders per step (BASELINE config C5: 64 views per step on 8 GPUs = 8 per GPU)")
    ap.add_argument("--config", default="C3")
    ap.add_argument("--gaussians", type=int, default=0, help="override P (debug only; makes the number invalid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--batched", action="store_true",
                    help="render each rank's views of a step in ONE batched call (GaussianRasterizerBatched, row f1)")
    ap.add_argument("--no-train-step", action="store_true",
                    help="skip the extra `train_step` object (a full iteration: activations, render, L1+D-SSIM, Adam; N=1 only)")
    ap.add_argument("--plain-autograd", action="store_true",
                    help="sum the views' gradients through autograd (.grad += per parameter) instead of in the backward kernels")
    ap.add_argument("--grad-sum", default="deferred", choices=["deferred", "kernels", "autograd"],
                    help="how a step's views are summed: deferred = blending backward per view, ONE per-Gaussian backward per step "
                         "(set_grad_mode('deferred') + flush_deferred); kernels = every view's backward adds into the buffer "
                         "(set_grad_mode overwrite / accumulate); autograd = plain .grad += (same as --plain-autograd)")
    ap.add_argument("--pipeline", type=int, default=1,
                    help="deferred steps: 1 = the forward of view i+1 is issued before the backward of view i, its front (preprocess, "
                         "depth order, tile split) on a high-priority stream of its own (set_front_stream / gsr_forward_front); 0 = one "
                         "view at a time on one stream")
    ap.add_argument("--prepare", type=int, default=1,
                    help="deferred steps: 1 = the preprocess stage of the step's views runs in one pass over the parameters "
                         "(prepare_views / gsr_preprocess_views) before the per-view calls; 0 = every call preprocesses its view")
    ap.add_argument("--front-streams", type=int, default=2,
                    help="--pipeline 1: the views' fronts alternate over this many high-priority streams (2: the fronts of consecutive "
                         "views may run next to each other - 1.962 against 1.975 ms per view with 1)")
    ap.add_argument("--views-of-rank", default="", help="diagnostic (N = 1): r/w = render the cameras rank r of w ranks would get")
    ap.add_argument("--nccl-priority", default="high", choices=["high", "default"], help="priority of NCCL's stream (N > 1)")
    ap.add_argument("--front-priority", default="high", choices=["high", "default"],
                    help="priority of the front stream of --pipeline 1 (A/B: `default` = same priority as the rendering stream)")
    ap.add_argument("--no-extras", action="store_true",
                    help="headline + stages only: skip batched / two-stream / forward-only / plain-api / dist2 / reference-cuda / c5 legs")
    ap.add_argument("--no-c5", action="store_true", help="skip the config C5 leg (6 M Gaussians)")
    ap.add_argument("--overlap-chunks", type=int, default=4,
                    help="N>1: ranges of the last view's per-Gaussian backward whose all-reduce is issued behind them (0 = one all-reduce)")
    ap.add_argument("--ref-kind", default="auto", choices=["auto", "cuda", "cpu"],
                    help="--impl reference: the reference CUDA extension on the GPU (auto when it is built and a GPU exists) or the CPU restatement")
    ap.add_argument("--nccl-debug", default="", help="write NCCL_DEBUG=INFO output (algorithm / protocol lines) to this file prefix")
    return ap.parse_args()


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock and throttle reasons with NVML every 50 ms while the timed region runs.  (Every 10 ms was too often:
    the headline pass - the only one the sampler runs next to - came out at 1.99-2.45 ms per view in some processes while every
    other pass of the same process measured the same step at 1.91; NVML queries take a driver lock kernel launches also need.)"""

    def __init__(self, index: int):
        self.index, self.samples, self.reasons, self._stop, self.max_mhz = index, [], set(), threading.Event(), None
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t is not None:
            self._t.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------ CPU baseline
def cpu_sample(spec, P, tile_stride=1, cam_k=0):
    """The oracle (CPU restatement of the reference algorithm, OpenMP over all host threads) on one view of the
    workload: preprocess + binning of the FULL scene, blending forward/backward on every `tile_stride`-th tile
    (1 = the whole image; > 1 scales the blending time by tile_stride and is flagged "extrapolated").
    Returns (seconds/view, info)."""
    import oracle
    from gsr_b200 import synth
    sc = synth.spec_scene(spec, P=P)
    cam = synth.make_camera(cam_k, 64, spec.W, spec.H, spec.radius, spec.fovx_deg)
    bg = synth.spec_background(spec)
    go = synth.grad_image(spec.W, spec.H).numpy()
    kw = dict(shs=sc["shs"].numpy(), scales=sc["scales"].numpy(), rotations=sc["rotations"].numpy(),
              viewmatrix=cam.viewmatrix.numpy(), projmatrix=cam.projmatrix.numpy(), campos=cam.campos.numpy(), bg=bg.numpy(),
              W=spec.W, H=spec.H, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, sh_degree=spec.D)
    m, o = sc["means3D"].numpy(), sc["opacities"].numpy()
    t0 = time.perf_counter()
    st = oracle.forward(m, o, blend=False, **kw)
    t_pre = time.perf_counter() - t0
    t0 = time.perf_counter()
    oracle.blend_forward(st, tile_first=0, tile_stride=tile_stride)
    t_bf = time.perf_counter() - t0
    t_pb = 0.0
    if tile_stride > 1:
        t0 = time.perf_counter()
        oracle.backward(st, go, tile_first=0, tile_stride=1 << 30)        # tile 0 only: the per-Gaussian part
        t_pb = time.perf_counter() - t0
    t0 = time.perf_counter()
    oracle.backward(st, go, tile_first=0, tile_stride=tile_stride)
    t_b = time.perf_counter() - t0
    if tile_stride > 1:
        est = t_pre + t_pb + tile_stride * (t_bf + max(0.0, t_b - t_pb))
    else:
        est = t_pre + t_bf + t_b
    info = dict(R=int(st.R), preprocess_bin_s=round(t_pre, 3), blend_fwd_s=round(t_bf, 3), backward_s=round(t_b, 3),
                tile_stride=tile_stride)
    return est, info


CPU_SAMPLE = "one full view of the workload per step (preprocess, binning, blending forward+backward over every tile, per-Gaussian backward)"


def cpu_baseline_object(spec, P, max_views=2, budget_s=30.0):
    """oracle/ C restatement on the box's host cores, bounded to ~10-30 s: `kind: port` (the reference has no CPU path)."""
    import oracle
    cores = oracle.set_num_threads(host_threads())        # launchers export OMP_NUM_THREADS=1: ask for every core we may use
    ests, info = [], None
    for k in range(max_views):
        est, info = cpu_sample(spec, P, 1, cam_k=k)
        ests.append(est)
        if sum(ests) > budget_s:
            break
    est = sum(ests) / len(ests)
    return {"value": 1.0 / est, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d full views of the workload (every tile), mean" % len(ests), "detail": info}


# ------------------------------------------------------------------------------------------ reference CUDA extension
def load_reference_extension():
    """(ref_dgr_C, ref_knn_C) of oracle/_ref - the unmodified reference extension built by oracle/build_ref.sh - or (None, None)."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(ref_dir) or not any(f.startswith("ref_dgr_C") for f in os.listdir(ref_dir)):
        return None, None
    if ref_dir not in sys.path:
        sys.path.insert(0, ref_dir)
    try:
        import torch  # noqa: F401
        import ref_dgr_C
        import ref_knn_C
        return ref_dgr_C, ref_knn_C
    except Exception:
        return None, None


def make_reference_function(refC):
    """The reference's autograd Function (diff_gaussian_rasterization/__init__.py:44-155: same argument and gradient order)
    over the reference's own _C module: plain autograd, no extension of ours anywhere on this path."""
    import torch

    class RefRasterize(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, rs):
            args = (rs["bg"], means3D, colors_precomp, opacities, scales, rotations, rs["scale_modifier"], cov3Ds_precomp,
                    rs["viewmatrix"], rs["projmatrix"], rs["tanfovx"], rs["tanfovy"], rs["H"], rs["W"], sh, rs["sh_degree"],
                    rs["campos"], False, False)
            num_rendered, color, radii, geom, binning, img = refC.rasterize_gaussians(*args)
            ctx.rs, ctx.num_rendered = rs, num_rendered
            ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img)
            return color, radii

        @staticmethod
        def backward(ctx, grad_out_color, _):
            rs = ctx.rs
            colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img = ctx.saved_tensors
            g = refC.rasterize_gaussians_backward(rs["bg"], means3D, radii, colors_precomp, scales, rotations, rs["scale_modifier"],
                                                  cov3Ds_precomp, rs["viewmatrix"], rs["projmatrix"], rs["tanfovx"], rs["tanfovy"],
                                                  grad_out_color, sh, rs["sh_degree"], rs["campos"], geom, ctx.num_rendered, binning,
                                                  img, False)
            g2d, gcol, gop, g3d, gcov, gsh, gsc, grot = g
            return g3d, g2d, gsh, None, gop, gsc, grot, None, None

    return RefRasterize


def reference_render(RefF, params, cam, bg, D):
    import torch
    e = torch.Tensor([])
    rs = dict(bg=bg, scale_modifier=1.0, viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix, tanfovx=cam.tanfovx,
              tanfovy=cam.tanfovy, H=cam.H, W=cam.W, sh_degree=D, campos=cam.campos)
    means2D = torch.zeros_like(params["means3D"], requires_grad=True)
    return RefF.apply(params["means3D"], means2D, params["shs"], e, params["opacities"], params["scales"], params["rotations"], e, rs)


def run_reference_arm(args, spec, P):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return                                           # rank 0 alone runs the reference arm; the others exit 0
    kind = args.ref_kind
    refC = None
    if kind in ("auto", "cuda"):
        try:
            import torch
            if torch.cuda.is_available():
                refC, _ = load_reference_extension()
        except Exception:
            refC = None
        if refC is None and kind == "cuda":
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref (reference CUDA extension) is not built or no GPU"}), flush=True)
            return
    if refC is not None:
        return run_reference_cuda(args, spec, P, refC)
    return run_reference_cpu(args, spec, P)


def run_reference_cuda(args, spec, P, refC):
    """The unmodified reference extension on this GPU: same scene, cameras, views per step; plain autograd accumulation."""
    import torch
    from gsr_b200 import synth
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    vps, V = args.views_per_step, 64
    scene = synth.spec_scene(spec, P=P)
    names = ("means3D", "scales", "rotations", "opacities", "shs")
    params = {k: scene[k].to(dev).clone().requires_grad_(True) for k in names}
    cams = [synth.make_camera(k, V, spec.W, spec.H, spec.radius, spec.fovx_deg).to(dev) for k in range(V)]
    bg = synth.spec_background(spec, dev)
    go = synth.grad_image(spec.W, spec.H, device=dev)
    RefF = make_reference_function(refC)

    def step(s):
        for p_ in params.values():
            p_.grad = None
        for j in range(vps):
            color, _radii = reference_render(RefF, params, cams[(s * vps + j) % V], bg, spec.D)
            color.backward(go)

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    value = args.steps * vps / (ms * 1e-3)
    cpu = None
    if not args.no_cpu_baseline:
        cpu = cpu_baseline_object(spec, P, max_views=1)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "ms_per_view": ms / args.steps / vps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "P": P, "W": spec.W, "H": spec.H, "sh_degree": spec.D, "views_per_step": vps,
                       "views_per_gpu_per_step": vps, "grad_sum": "autograd .grad +=",
                       "api": "the reference's _RasterizeGaussians over its own _C module, one call per view",
                       "device": "one GPU (the reference has no multi-GPU path: rank 0 alone runs this arm)"},
            "reference_kind": "reference CUDA extension (oracle/_ref, unmodified sources of /root/reference built by oracle/build_ref.sh) "
                              "on the same GPU; the reference has no CPU implementation",
            "cpu_baseline": cpu,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "clocks": clocks}
    print(json.dumps(line), flush=True)


def run_reference_cpu(args, spec, P):
    import oracle
    cores = oracle.set_num_threads(host_threads())
    for _ in range(min(args.warmup, 1)):
        cpu_sample(spec, P, 1 << 12)                 # warm-up on a tiny tile sample (page-in, OpenMP pool)
    t = []
    info = None
    for i in range(args.steps):
        est, info = cpu_sample(spec, P, 1, cam_k=i % 64)
        t.append(est)
        if sum(t) > 150.0:                           # keep the whole arm within a few minutes on slow hosts
            break
    sec_per_view = sum(t) / len(t)
    value = 1.0 / sec_per_view
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": len(t),
            "warmup": args.warmup, "ms_per_step": sec_per_view * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "P": P, "W": spec.W, "H": spec.H, "sh_degree": spec.D, "views_per_step": 1},
            "reference_kind": "CPU restatement of the reference algorithm (oracle/): no GPU or no built reference extension here",
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": CPU_SAMPLE, "detail": info},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "the reference rasterizer has no CPU implementation; this is the oracle/ C restatement of its algorithm"}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ our arm
def stage_bytes(P, Pv, R, N, tiles, M, accumulate_frac=0.0):
    """Algorithmic bytes per launch of each stage."""
    return {
        # inputs (44 B) + SH of the visible ones + 48 B record, clamp byte, radii, depth key, tile rect
        "preprocess": 44 * P + (12 * M + 4) * Pv + (48 + 1 + 4 + 4 + 8) * P,
        # digit histograms (4P) + four passes over (key, id) pairs (first reads keys only, last writes ids only) + tile rects (8P)
        "depth_order": 4 * P + (4 * 16 - 8) * P + 8 * P + 8 * tiles,
        # emission: order + rect in, (tile, id) out; pass 1: 8R in, 8R out; pass 2: 8R in, 4R out
        "tile_split": 12 * P + 8 * R + 16 * R + 12 * R,
        "blend_forward": 40 * R + 20 * N,
        "blend_backward": 80 * R + 32 * N,
        # every owned gradient written once (zeros for culled rows) + inputs; accumulating also reads the old gradients
        "preprocess_backward": (111 + 64 + 24 * M) * Pv + 4 * (14 + 3 * M) * (P - Pv) * (1.0 - accumulate_frac)
                               + accumulate_frac * 4 * (11 + 3 * M) * Pv,
    }


def main():
    args = parse_args()
    from gsr_b200 import synth
    spec = synth.CONFIGS[args.config]
    P = args.gaussians or spec.P
    if args.impl == "reference":
        run_reference_arm(args, spec, P)
        return

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.nccl_debug and world > 1:
        os.environ["NCCL_DEBUG"] = "INFO"
        os.environ["NCCL_DEBUG_SUBSYS"] = "INIT,ENV,TUNING"
        os.environ["NCCL_DEBUG_FILE"] = args.nccl_debug + ".rank%d.log" % rank
    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL's kernels on a high-priority stream: a range's all-reduce is issued while the next range of the per-Gaussian
        # backward fills every SM, and an NCCL CTA needs most of an SM to itself - at default priority it waits for the tail
        from gsr_b200 import dp as _dp
        _dp.init_nccl(dev, high_priority=args.nccl_priority == "high")
    import diff_gaussian_rasterization as dgr
    from gsr_b200 import dp, losses
    C = dgr._C
    extras = not args.no_extras

    V = 64
    vps = args.views_per_step
    scene = synth.spec_scene(spec, P=P)
    params = dp.make_leaf_params(scene, dev)
    flat = dp.FlatGrads(params, nccl_registered=world > 1)
    cams = [synth.make_camera(k, V, spec.W, spec.H, spec.radius, spec.fovx_deg).to(dev) for k in range(V)]
    bg = synth.spec_background(spec, dev)
    go = synth.grad_image(spec.W, spec.H, device=dev)
    M = (spec.D + 1) ** 2
    N = spec.W * spec.H
    tiles = ((spec.W + 15) // 16) * ((spec.H + 15) // 16)

    vr, vw = rank, world
    if args.views_of_rank:
        # (diagnostic, N = 1: render the camera set rank r of w would get - how uneven are the ranks' loads?)
        vr, vw = (int(x) for x in args.views_of_rank.split("/"))

    def my_views(step):
        # view k of the step goes to rank k mod world; camera ids walk round the ring
        return [(step * vps * vw + j * vw + vr) % V for j in range(vps)]

    stats = {"R": 0, "Pv": 0, "views": 0}
    if args.plain_autograd:
        args.grad_sum = "autograd"
    fused = args.grad_sum != "autograd"
    deferred = args.grad_sum == "deferred" and not args.batched
    overlap = args.overlap_chunks if (world > 1 and fused and not args.batched) else 0

    def grad_mode(i):
        # view 0 of the step overwrites the flat gradient buffer, the following views add to it (in the kernels)
        return dgr.set_grad_mode(("overwrite" if i == 0 else "accumulate") if fused else None)

    from diff_gaussian_rasterization.batched import BatchedRasterizationSettings, GaussianRasterizerBatched
    vm_all = torch.stack([c.viewmatrix for c in cams]); pm_all = torch.stack([c.projmatrix for c in cams])
    cp_all = torch.stack([c.campos for c in cams])
    go_b = go.unsqueeze(0).expand(vps, -1, -1, -1).contiguous()

    def render_batched(prm, ks, vm, pm, cp, npts):
        rs = BatchedRasterizationSettings(spec.H, spec.W, cams[0].tanfovx, cams[0].tanfovy, bg, 1.0, vm, pm, spec.D, cp, False)
        means2D = torch.zeros(len(ks), npts, 3, device=dev, requires_grad=True)
        return GaussianRasterizerBatched(rs)(prm["means3D"], means2D, prm["opacities"], prm["shs"], prm["scales"], prm["rotations"])

    def step_device_batched(step, collect=False):
        ks = torch.tensor(my_views(step), device=dev)
        prev = dgr.set_grad_mode("overwrite")
        color, radii = render_batched(params, ks, vm_all[ks], pm_all[ks], cp_all[ks], P)
        color.backward(go_b)
        dgr.set_grad_mode(prev)
        if collect:
            stats["Pv"] += int((radii > 0).sum())
            stats["views"] += len(ks)
        flat.all_reduce()

    use_deferred = [True]            # (switched off for the `grad_sum_in_kernels` comparison leg)
    front_st = lambda j=0: dgr.front_stream(dev, args.front_priority == "high", j)
    n_fronts = max(1, args.front_streams)

    def fronts_begin():
        # every front stream is ordered behind what is queued so far (the parameters); the first one stays set
        prev_ = dgr.set_front_stream(front_st(n_fronts - 1))
        for j in range(n_fronts - 2, -1, -1):
            dgr.set_front_stream(front_st(j))
        return prev_

    def front_for(i):
        # view i's front goes to stream i mod n (no further wait: the streams were ordered at the start of the step)
        if n_fronts > 1:
            dgr.set_front_stream(front_st(i % n_fronts), wait=False)
    use_prepare = [bool(args.prepare) and deferred]        # (switched off with the pipelining for the `one_stream` leg)
    use_pipeline = [bool(args.pipeline) and deferred]      # (switched off for the per-stage timing pass and the `one_stream` leg)

    def run_views(prm, flt, views, loss_of, chunks, collect=False):
        """One rank's share of a step through the drop-in API: the views' gradients are summed in the backward kernels; with
        chunks > 0 the LAST view's per-Gaussian backward runs in ranges and each range's all-reduce starts behind it."""
        works = []
        if deferred and use_deferred[0]:
            # per view: forward + blending backward; once per step: the per-Gaussian backward of all views (in `chunks` ranges
            # with each range's all-reduce issued behind it when N > 1)
            views = list(views)
            piped = use_pipeline[0] and len(views) > 1
            prev = dgr.set_grad_mode("deferred")
            prev_front = fronts_begin() if piped else None

            def back(color, radii, k):
                loss_of(color, k).backward() if loss_of is not None else color.backward(go)
                if collect:
                    stats["Pv"] += int((radii > 0).sum())
                    stats["views"] += 1
            try:
                if use_prepare[0]:
                    dp.prepare_step_views(dgr, prm, cams, bg, spec.D, views)
                if piped:
                    # software pipeline: forward of view i+1 (its front on the high-priority stream), then backward of view i
                    front_for(0)
                    ahead = dp.render_view(dgr, prm, cams[views[0]], bg, spec.D)
                    for i, k in enumerate(views):
                        color, radii, _ = ahead
                        front_for(i + 1)
                        ahead = dp.render_view(dgr, prm, cams[views[i + 1]], bg, spec.D) if i + 1 < len(views) else None
                        back(color, radii, k)
                else:
                    for k in views:
                        color, radii, _ = dp.render_view(dgr, prm, cams[k], bg, spec.D)
                        back(color, radii, k)
            finally:
                dgr.set_grad_mode(prev)
                if piped:
                    dgr.set_front_stream(prev_front)
            if chunks > 0:
                dgr.flush_deferred(False, chunks, lambda g0, g1: works.append(flt.all_reduce_rows(g0, g1)))
                for w in works:
                    if w is not None:
                        w.wait()
            else:
                dgr.flush_deferred(False)
                flt.all_reduce()
            return
        for i, k in enumerate(views):
            prev = grad_mode(i)
            hook = None
            last = chunks > 0 and i == len(views) - 1
            if last:
                hook = dgr.set_chunk_hook((chunks, lambda g0, g1: works.append(flt.all_reduce_rows(g0, g1))))
            try:
                color, radii, _ = dp.render_view(dgr, prm, cams[k], bg, spec.D)
                loss_of(color, k).backward() if loss_of is not None else color.backward(go)
            finally:
                dgr.set_grad_mode(prev)
                if last:
                    dgr.set_chunk_hook(hook)
            if collect:
                stats["Pv"] += int((radii > 0).sum())
                stats["views"] += 1
        if chunks > 0 and works:
            for w in works:
                if w is not None:
                    w.wait()
        else:
            flt.all_reduce()

    def step_device(step, collect=False):
        if args.batched:
            return step_device_batched(step, collect)
        if not fused:
            flat.zero_()
        run_views(params, flat, my_views(step), None, overlap, collect)

    def step_device_streams(step):
        # same step through gsr_b200.dp.backward_views with the views alternating over two CUDA streams (two views in flight)
        dp.backward_views(dgr, params, flat, cams, bg, spec.D, lambda color, k: (color * go).sum(), my_views(step), fused=True,
                          n_streams=2)
        flat.all_reduce()

    # e2e: host buffers in, host scalar out.  Per view: camera + ground-truth image copied from pinned host memory
    # (on a copy stream, one view ahead of the rendering, as a trainer's data loader does), forward, fused L1 loss
    # (gsr_b200.losses) against it, backward; the step's summed loss is read back to the host once per step (train.py:91-106).
    target_host = torch.rand(3, spec.H, spec.W, generator=torch.Generator().manual_seed(3)).pin_memory()
    cam_host = [torch.cat([c.viewmatrix.reshape(-1), c.projmatrix.reshape(-1), c.campos.reshape(-1)]).cpu().pin_memory()
                for c in cams]
    h2d_view = target_host.numel() * 4 + cam_host[0].numel() * 4
    copy_stream = torch.cuda.Stream(device=dev)
    pending = {}
    # landing buffers for the prefetched views: a ring of three, allocated once, so the timed region makes no allocator calls
    # (a cudaMalloc on the copy stream stalls the whole device); a slot is overwritten only after the compute stream is
    # done with the view that used it
    RING = 4
    ring_tgt = [torch.empty(3, spec.H, spec.W, device=dev) for _ in range(RING)]
    ring_cam = [torch.empty(cam_host[0].numel(), device=dev) for _ in range(RING)]
    ring_free = [None] * RING
    ring_next = [0]

    def prefetch(k):
        slot = ring_next[0] % RING
        ring_next[0] += 1
        with torch.cuda.stream(copy_stream):
            if ring_free[slot] is not None:
                copy_stream.wait_event(ring_free[slot])
            ring_tgt[slot].copy_(target_host, non_blocking=True)
            ring_cam[slot].copy_(cam_host[k], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        pending[k] = (slot, ev)

    # cameras of a whole step (prepared steps): [views, 35] floats, a ring of three buffers
    cam_ring = [torch.empty(vps, cam_host[0].numel(), device=dev) for _ in range(3)]
    cam_pending = {}

    def prefetch_cams(step):
        buf = cam_ring[step % 3]
        with torch.cuda.stream(copy_stream):
            # (the buffer's previous user is two steps back: its renders were queued before the loss read-back of that step)
            for j, k in enumerate(my_views(step)):
                buf[j].copy_(cam_host[k], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        cam_pending[step] = (buf, ev)

    def take_cams(step):
        if step not in cam_pending:
            prefetch_cams(step)
        buf, ev = cam_pending.pop(step)
        torch.cuda.current_stream().wait_event(ev)
        if use_pipeline[0]:
            for j in range(n_fronts):
                front_st(j).wait_event(ev)
        return buf

    def take(k):
        if k not in pending:
            prefetch(k)
        slot, ev = pending.pop(k)
        torch.cuda.current_stream().wait_event(ev)
        if use_pipeline[0]:
            for j in range(n_fronts):
                front_st(j).wait_event(ev)   # the camera is an input of the front too
        return ring_tgt[slot], ring_cam[slot], slot

    def release(slot):
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        ring_free[slot] = ev

    def prefetch_step(step):
        with torch.cuda.stream(copy_stream):
            ks = my_views(step)
            tgt = torch.empty(len(ks), 3, spec.H, spec.W, device=dev)
            cm = torch.empty(len(ks), cam_host[0].numel(), device=dev)
            for j, k in enumerate(ks):
                tgt[j].copy_(target_host, non_blocking=True)
                cm[j].copy_(cam_host[k], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        pending[("step", step)] = (tgt, cm, ev)

    def step_e2e_batched(step):
        if ("step", step) not in pending:
            prefetch_step(step)
        tgt, cmat, ev = pending.pop(("step", step))
        cur = torch.cuda.current_stream()
        cur.wait_event(ev)
        tgt.record_stream(cur); cmat.record_stream(cur)
        prefetch_step(step + 1)                      # next step's inputs travel while this one renders
        prev = dgr.set_grad_mode("overwrite")
        color, radii = render_batched(params, my_views(step), cmat[:, :16].reshape(-1, 4, 4), cmat[:, 16:32].reshape(-1, 4, 4),
                                      cmat[:, 32:35].contiguous(), P)
        loss = losses.l1_loss(color, tgt)
        loss.backward()
        dgr.set_grad_mode(prev)
        flat.all_reduce()
        _ = float(loss.item())

    def step_e2e(step):
        if args.batched:
            return step_e2e_batched(step)
        if not fused:
            flat.zero_()
        views = my_views(step)
        total = None
        works = []
        if deferred:
            piped = use_pipeline[0] and len(views) > 1
            prev = dgr.set_grad_mode("deferred")
            prev_front = fronts_begin() if piped else None
            step_cams = None
            if use_prepare[0]:
                # the step's cameras (a trainer samples them before the step): one small copy per view from pinned host
                # memory on the copy stream, a step ahead; the views' preprocess then runs in one pass over the parameters
                cmb = take_cams(step)
                prefetch_cams(step + 1)
                step_cams = [synth.Camera(spec.W, spec.H, cams[k].tanfovx, cams[k].tanfovy, cmb[j, :16].view(4, 4),
                                          cmb[j, 16:32].view(4, 4), cmb[j, 32:35]) for j, k in enumerate(views)]
                dp.prepare_step_views(dgr, params, step_cams, bg, spec.D, range(len(views)))

            def fwd(i):
                k = views[i]
                if piped:
                    front_for(i)
                tgt, cm, slot = take(k)
                nxt = views[i + 1] if i + 1 < len(views) else my_views(step + 1)[0]
                prefetch(nxt)                              # next view's inputs travel while this one renders
                cam = step_cams[i] if step_cams is not None else synth.Camera(
                    spec.W, spec.H, cams[k].tanfovx, cams[k].tanfovy, cm[:16].view(4, 4), cm[16:32].view(4, 4), cm[32:35])
                color, radii, _ = dp.render_view(dgr, params, cam, bg, spec.D)
                return color, tgt, slot

            try:
                ahead = fwd(0) if piped else None
                for i, k in enumerate(views):
                    if piped:
                        color, tgt, slot = ahead
                        ahead = fwd(i + 1) if i + 1 < len(views) else None
                    else:
                        color, tgt, slot = fwd(i)
                    loss = losses.l1_loss(color, tgt)
                    loss.backward()
                    release(slot)
                    total = loss.detach() if total is None else total + loss.detach()
            finally:
                dgr.set_grad_mode(prev)
                if piped:
                    dgr.set_front_stream(prev_front)
            if overlap > 0:
                dgr.flush_deferred(False, overlap, lambda g0, g1: works.append(flat.all_reduce_rows(g0, g1)))
                for w in works:
                    if w is not None:
                        w.wait()
            else:
                dgr.flush_deferred(False)
                flat.all_reduce()
            _ = float(total.item())
            return
        for i, k in enumerate(views):
            prev = grad_mode(i)
            last = overlap > 0 and i == len(views) - 1
            hook = dgr.set_chunk_hook((overlap, lambda g0, g1: works.append(flat.all_reduce_rows(g0, g1)))) if last else None
            try:
                tgt, cm, slot = take(k)
                nxt = views[i + 1] if i + 1 < len(views) else my_views(step + 1)[0]
                prefetch(nxt)                                  # next view's inputs travel while this one renders
                cam = synth.Camera(spec.W, spec.H, cams[k].tanfovx, cams[k].tanfovy, cm[:16].view(4, 4), cm[16:32].view(4, 4), cm[32:35])
                color, radii, _ = dp.render_view(dgr, params, cam, bg, spec.D)
                loss = losses.l1_loss(color, tgt)
                loss.backward()
                release(slot)
            finally:
                dgr.set_grad_mode(prev)
                if last:
                    dgr.set_chunk_hook(hook)
            total = loss.detach() if total is None else total + loss.detach()
        if works:
            for w in works:
                if w is not None:
                    w.wait()
        else:
            flat.all_reduce()
        _ = float(total.item())             # device -> host read of the step's result

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, first_step):
        # no collector pause of the interpreter inside a timed region: the host runs only two views ahead of the GPU, so a
        # full collection (10-15 ms with this many live objects) drains the queues - the 1.99-2.03 ms runs among the 1.91 ones
        import gc
        gc.collect()
        gc_was = gc.isenabled()
        gc.disable()
        barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(first_step + i)
        e1.record()
        barrier()
        if gc_was:
            gc.enable()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- device-resident timing (value) with per-stage events and clock sampling
    for i in range(args.warmup):
        step_device(i)
    torch.cuda.synchronize()
    step_device(args.warmup, collect=True)           # untimed: realised P_v for the byte counts
    piped = use_pipeline[0] and not args.batched and vps > 1
    C.timing_reset(); C.timing_enable(not piped); C.reset_launch_count()
    sampler = ClockSampler(local_rank)
    sampler.start()
    dev_allocs0 = torch.cuda.memory_stats(dev).get("num_device_alloc", 0)
    ms_total = timed(step_device, args.steps, args.warmup + 1)
    dev_allocs = torch.cuda.memory_stats(dev).get("num_device_alloc", 0) - dev_allocs0     # cudaMallocs inside the timed region
    clocks = sampler.stop()
    C.timing_enable(False)
    launches = int(C.launch_count())
    one_stream = None
    if piped:
        # The per-stage event intervals overlap when two views are in flight, so the stage table (and the dominant kernel's
        # launch duration the roofline uses) comes from a pass of the same steps WITHOUT the pipelining, timed as a whole too.
        use_pipeline[0] = False
        for w_ in range(2):                 # (the allocator settles on this form's block sizes)
            step_device(args.warmup + w_)
        C.timing_reset(); C.timing_enable(True)
        ms_one = timed(step_device, args.steps, args.warmup + 1)
        C.timing_enable(False)
        use_pipeline[0] = True
        one_stream = {"api": "the same step, one view at a time on one stream (--pipeline 0); the `stages` table is from this pass",
                      "ms_per_view": ms_one / args.steps / vps, "value": args.steps * vps * world / (ms_one * 1e-3), "unit": UNIT}
        if use_prepare[0]:
            # ... and with every call preprocessing its own view as well (round 2 g's form of the step)
            use_pipeline[0] = False; use_prepare[0] = False
            for w_ in range(2):
                step_device(args.warmup + w_)
            ms_pv = timed(step_device, args.steps, args.warmup + 1)
            use_pipeline[0] = True; use_prepare[0] = True
            one_stream["per_view_preprocess_ms_per_view"] = ms_pv / args.steps / vps
    stage_rows = C.timing_collect()
    stages_piped = None
    if piped:
        # the same stage intervals with two views in flight: how far each stage is stretched by what runs next to it
        C.timing_reset(); C.timing_enable(True)
        timed(step_device, args.steps, args.warmup + 1)
        C.timing_enable(False)
        per_step = ("preprocess_backward", "preprocess") if use_prepare[0] else ("preprocess_backward",)
        stages_piped = {name: round(ms / (args.steps * vps if name in per_step else max(1, calls)), 4)
                        for name, ms, calls in C.timing_collect() if calls}
    views_total = args.steps * vps * world
    value = views_total / (ms_total * 1e-3)

    # ---- e2e timing
    e2e = None
    if not args.no_e2e:
        for i in range(max(3, args.warmup)):      # untimed: allocator, pinned buffers and NCCL settle
            step_e2e(i)
        ms_e2e = timed(step_e2e, args.steps, args.warmup)
        e2e = {"value": views_total / (ms_e2e * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": h2d_view * vps + (cam_host[0].numel() * 4 * vps if (use_prepare[0] and not args.batched) else 0),
               "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / args.steps,
               "note": "per view: 24.9 MB target image + camera from pinned host memory on a copy stream one view ahead (ring of "
                       "four preallocated landing buffers); loss read once per step"}

    # ---- the collective by itself (N > 1): the flat buffer's all-reduce timed alone, and what the step pays for it
    allreduce = None
    if world > 1:
        for _ in range(2):
            flat.all_reduce()
        barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for _ in range(reps):
            flat.all_reduce()
        e1.record()
        barrier()
        t_ar = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
        dist.all_reduce(t_ar, op=dist.ReduceOp.MAX)
        t_ar = float(t_ar.item())
        allreduce = {"bytes": flat.nbytes(), "ms_alone": round(t_ar, 4),
                     "busbw_GBps": round(2.0 * (world - 1) / world * flat.nbytes() / (t_ar * 1e-3) / 1e9, 1),
                     "buffer": "NCCL-allocated and registered (ncclMemAlloc / ncclCommRegister via torch.cuda.MemPool)" if flat.registered
                               else "plain device buffer (registration unavailable: %s)" % getattr(flat, "register_error", "n/a"),
                     "overlap_chunks": overlap,
                     "limits": "the fp32 gradient all-reduce (236 B per Gaussian at SH-3) after the rank's last view; only the last view's "
                               "per-Gaussian backward (~0.4 ms) can hide part of it, plus the wait for the slowest rank"}

    # ---- the same step through the batched entry point (one call per rank and step), reported next to the drop-in API's
    batched_extra = None
    if extras and not args.batched:
        for i in range(args.warmup):
            step_device_batched(i)
        pending.clear()
        ms_b = timed(step_device_batched, args.steps, args.warmup)
        batched_extra = {"api": "GaussianRasterizerBatched (gsr_forward_batched): one call per rank and step",
                         "value": views_total / (ms_b * 1e-3), "unit": UNIT, "ms_per_view": ms_b / args.steps / vps}
        if not args.no_e2e:
            for i in range(max(3, args.warmup)):      # untimed: allocator, pinned buffers and NCCL settle
                step_e2e_batched(i)
            ms_be = timed(step_e2e_batched, args.steps, args.warmup)
            batched_extra["e2e"] = {"value": views_total / (ms_be * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d_view * vps,
                                    "d2h_bytes_per_step": 4}
        pending.clear()

    # ---- config C2: forward only under torch.no_grad(), same scene and cameras
    def step_forward_only(step):
        with torch.no_grad():
            for k in my_views(step):
                dp.render_view(dgr, params, cams[k], bg, spec.D)

    forward_extra = None
    if extras and not args.batched:
        for i in range(2):
            step_forward_only(i)
        ms_f = timed(step_forward_only, args.steps, args.warmup)
        forward_extra = {"config": "C2: forward only, torch.no_grad()", "ms_per_view": ms_f / args.steps / vps,
                         "value": views_total / (ms_f * 1e-3), "unit": UNIT}

    # ---- the same step with two views in flight (two CUDA streams through the drop-in API), reported next to the headline
    streams_extra = None
    if extras and not args.batched and fused and vps > 1:
        for i in range(args.warmup):
            step_device_streams(i)
        ms_s = timed(step_device_streams, args.steps, args.warmup)
        streams_extra = {"api": "GaussianRasterizer from two CUDA streams (gsr_b200.dp.backward_views(n_streams=2))",
                         "value": views_total / (ms_s * 1e-3), "unit": UNIT, "ms_per_view": ms_s / args.steps / vps}

    # ---- the same step with every view's backward adding into the buffer itself (round 1's headline form), next to the deferred one
    kernels_extra = None
    if extras and deferred:
        use_deferred[0] = False
        for i in range(args.warmup):
            step_device(i)
        ms_k = timed(step_device, args.steps, args.warmup)
        use_deferred[0] = True
        kernels_extra = {"api": "GaussianRasterizer, set_grad_mode overwrite / accumulate: the per-Gaussian backward runs per view",
                         "value": views_total / (ms_k * 1e-3), "unit": UNIT, "ms_per_view": ms_k / args.steps / vps}

    # ---- what an UNCHANGED train.py sees (gaussian_renderer/__init__.py:53-80): raw parameters, torch activations,
    # torch.cat of the two SH parameters, plain autograd (no set_grad_mode, no split storage): every view pays the
    # activations, the concatenation and the elementwise `.grad +=`
    plain_extra = None
    if extras and world == 1 and not args.batched:
        leaf = lambda t: t.to(dev).clone().contiguous().requires_grad_(True)
        raw = {"xyz": leaf(scene["means3D"]), "f_dc": leaf(scene["shs"][:, :1]), "f_rest": leaf(scene["shs"][:, 1:]),
               "scaling": leaf(torch.log(scene["scales"])), "rotation": leaf(scene["rotations"]),
               "opacity": leaf(torch.logit(scene["opacities"].clamp(1e-4, 1 - 1e-4)))}

        def step_plain(step):
            for t in raw.values():
                t.grad = None
            for k in my_views(step):
                prm = {"means3D": raw["xyz"], "shs": torch.cat((raw["f_dc"], raw["f_rest"]), dim=1), "scales": torch.exp(raw["scaling"]),
                       "rotations": torch.nn.functional.normalize(raw["rotation"]), "opacities": torch.sigmoid(raw["opacity"])}
                color, _, _ = dp.render_view(dgr, prm, cams[k], bg, spec.D)
                color.backward(go)

        prev = dgr.set_grad_mode(None)
        for i in range(2):
            step_plain(i)
        ms_p = timed(step_plain, max(3, args.steps // 4), args.warmup)
        dgr.set_grad_mode(prev)
        plain_extra = {"api": "GaussianRasterizer exactly as gaussian_renderer.render() calls it: activations + torch.cat'd SH inputs, "
                              "plain autograd accumulation (no extension of ours)",
                       "ms_per_view": ms_p / max(3, args.steps // 4) / vps, "value": vps * 1e3 / (ms_p / max(3, args.steps // 4)), "unit": UNIT}
        del raw
        torch.cuda.empty_cache()

    # ---- the reference CUDA extension on this GPU, same view (the optimisation bar), and distCUDA2 (config C3
    # names its scale initialisation): baseline legs like cpu_baseline, never on the product path
    ref_extra, dist2_extra = None, None
    if extras and world == 1:
        refC, refK = load_reference_extension()
        e = torch.Tensor([])

        def ev_time(fn, n=5, warm=2):
            for _ in range(warm):
                fn()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(n):
                fn()
            b.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b) / n

        pts = params["means3D"].detach()
        import simple_knn._C as knn
        t_ours = ev_time(lambda: knn.distCUDA2(pts), n=3, warm=1)
        dist2_extra = {"P": P, "ours_ms": round(t_ours, 3), "ours_points_per_s": P / (t_ours * 1e-3)}
        if refC is not None:
            cam0 = cams[my_views(0)[0]]
            det = {k: v.detach() for k, v in params.items()}
            fargs = (bg, det["means3D"], e, det["opacities"], det["scales"], det["rotations"], 1.0, e, cam0.viewmatrix, cam0.projmatrix,
                     cam0.tanfovx, cam0.tanfovy, cam0.H, cam0.W, det["shs"], spec.D, cam0.campos, False, False)
            t_f = ev_time(lambda: refC.rasterize_gaussians(*fargs))
            Rr, _c, rad, geom, binning, img = refC.rasterize_gaussians(*fargs)
            t_b = ev_time(lambda: refC.rasterize_gaussians_backward(bg, det["means3D"], rad, e, det["scales"], det["rotations"], 1.0, e,
                                                                     cam0.viewmatrix, cam0.projmatrix, cam0.tanfovx, cam0.tanfovy, go,
                                                                     det["shs"], spec.D, cam0.campos, geom, Rr, binning, img, False))
            ref_extra = {"what": "unmodified reference extension (oracle/_ref) on this GPU, raw _C calls on one C3 view",
                         "forward_ms": round(t_f, 3), "backward_ms": round(t_b, 3), "ms_per_view": round(t_f + t_b, 3),
                         "speedup_of_headline": round((t_f + t_b) / (ms_total / args.steps / vps), 2)}
            del geom, binning, img
            t_ref = ev_time(lambda: refK.distCUDA2(pts), n=2, warm=1)
            dist2_extra.update({"reference_ms": round(t_ref, 3), "reference_points_per_s": P / (t_ref * 1e-3),
                                "speedup": round(t_ref / t_ours, 2)})
            del det, fargs, rad, _c
        del pts
        torch.cuda.empty_cache()

    # ---- a whole training iteration (train.py:95-140) assembled from the fused pieces, next to the torch formulation over the
    # same rasterizer (gsr_b200.train_step); one view per iteration like the reference trainer.  Extra information only.
    train_extra = None
    if extras and world == 1 and not args.no_train_step and spec.D == 3:
        from gsr_b200 import train_step
        gen = torch.Generator().manual_seed(3)
        gts = [torch.rand(3, spec.H, spec.W, generator=gen).to(dev) for _ in range(2)]
        train_extra = {"what": "activations + render + L1/D-SSIM loss + backward + densification stats + Adam, one view per iteration"}
        for key, is_fused, n_it in (("fused", True, 10), ("torch_formulation", False, 5)):
            leaf = lambda t: t.to(dev).clone().contiguous().requires_grad_(True)
            model = {"xyz": leaf(scene["means3D"]), "features_dc": leaf(scene["shs"][:, :1]), "features_rest": leaf(scene["shs"][:, 1:]),
                     "scaling": leaf(torch.log(scene["scales"])), "rotation": leaf(scene["rotations"]),
                     "opacity": leaf(torch.logit(scene["opacities"].clamp(1e-4, 1 - 1e-4)))}
            st = train_step.new_stats(P, dev)
            opt = train_step.make_optimizer(model, is_fused)
            fn = train_step.fused_step if is_fused else train_step.reference_step
            for k in range(3):
                fn(model, st, opt, cams[k % V], bg, gts[k % 2], spec.D)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record()
            for k in range(n_it):
                fn(model, st, opt, cams[(3 + k) % V], bg, gts[k % 2], spec.D)
            e1.record(); torch.cuda.synchronize()
            train_extra[key + "_ms_per_iteration"] = e0.elapsed_time(e1) / n_it
            del model, st, opt
            torch.cuda.empty_cache()

    # ---- realised sizes of one representative view and the FP32-roofline probe (before the C5 leg replaces the scene)
    R = probe_R(C, params, cams[my_views(0)[0]], bg, spec.D)
    pairs = None
    if rank == 0:
        try:
            pairs = probe_pairs(C, params, cams[my_views(0)[0]], bg, spec.D)
        except Exception as ex:                      # the probe is extra information: never fail the bench line over it
            pairs = str(ex)[:200]

    # ---- config C5 (BASELINE.json config 5): 6 M Gaussians, 8 views per GPU and step (64 per step on 8 GPUs), same step
    c5 = None
    if extras and not args.no_c5 and not args.batched and spec.name in ("C3", "C2") and not args.gaussians:
        spec5 = synth.CONFIGS["C5"]
        n_flat = flat.nbytes()
        flat.release()
        del params, flat, scene
        torch.cuda.empty_cache()
        scene5 = synth.spec_scene(spec5)
        params5 = dp.make_leaf_params(scene5, dev)
        flat5 = dp.FlatGrads(params5, nccl_registered=world > 1)
        del scene5
        steps5 = max(3, args.steps // 4)

        def step_c5(step):
            run_views(params5, flat5, my_views(step), None, overlap)

        for i in range(4):                           # (the allocator's pool of stream-shared blobs settles within three steps)
            step_c5(i)
        ms5 = timed(step_c5, steps5, 4)
        R5 = probe_R(C, params5, cams[my_views(0)[0]], bg, spec5.D)
        c5 = {"workload": "C5: 6M Gaussians, SH degree 3, 1920x1080, %d views per step (%d per GPU), view-sharded over %d GPU(s), fwd+bwd via "
                          "GaussianRasterizer API + gradient all-reduce" % (vps * world, vps, world),
              "value": steps5 * vps * world / (ms5 * 1e-3), "unit": UNIT, "ms_per_step": ms5 / steps5, "ms_per_view_per_gpu": ms5 / steps5 / vps,
              "steps": steps5, "R_per_view": R5, "allreduce_bytes": flat5.nbytes(), "registered_buffer": bool(flat5.registered)}
        if world > 1:
            for _ in range(2):
                flat5.all_reduce()
            barrier()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3):
                flat5.all_reduce()
            e1.record()
            barrier()
            t5 = torch.tensor([e0.elapsed_time(e1) / 3], device=dev)
            dist.all_reduce(t5, op=dist.ReduceOp.MAX)
            c5["allreduce_ms_alone"] = round(float(t5.item()), 4)
            c5["busbw_GBps"] = round(2.0 * (world - 1) / world * flat5.nbytes() / (float(t5.item()) * 1e-3) / 1e9, 1)
        flat_nbytes = n_flat
        flat5.release()
        del params5, flat5
        torch.cuda.empty_cache()
    else:
        flat_nbytes = flat.nbytes()
        flat.release()                               # (a NCCL-registered buffer must go before the process group does)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback")
        Pv = stats["Pv"] / max(1, stats["views"])
        sb = stage_bytes(P, Pv, R, N, tiles, M, accumulate_frac=(vps - 1) / vps if (fused and not deferred) else 0.0)
        if deferred:
            # per view: record, accumulator row, radii, clamp byte in, dL/dmean2D out (113 B per Gaussian); per step: parameters
            # in and gradients out once (44 + 12 M and 4 (11 + 3 M) B per Gaussian)
            sb["preprocess_backward"] = 113 * P + (44 + 12 * M + 4 * (11 + 3 * M)) * P / vps
            if use_prepare[0]:
                # prepared steps: parameters (44 B + the SH row) in once per step, every view's record / clamp byte / radii /
                # depth key / tile rect out
                sb["preprocess"] = (44 + 12 * M) * P / vps + (48 + 1 + 4 + 4 + 8) * P
        stages = {}
        dom, dom_ms = None, -1.0
        for name, ms, calls in stage_rows:
            if calls == 0 or name not in sb:
                continue
            # per view (a batched call covers vps views; so does the deferred per-Gaussian backward, once per step)
            if deferred and (name == "preprocess_backward" or (name == "preprocess" and use_prepare[0])):
                avg = ms / args.steps / vps                   # (calls = steps x ranges when the flush runs in ranges)
            else:
                avg = ms / calls / (vps if args.batched else 1)
            gbs = sb[name] / (avg * 1e-3) / 1e9
            stages[name] = {"ms": round(avg, 4), "algo_bytes": int(sb[name]), "GBps": round(gbs, 1), "frac_hbm": round(gbs / peak, 4)}
            if avg > dom_ms:
                dom, dom_ms = name, avg
        prof = {}
        try:
            prof = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        except Exception:
            pass
        traffic = prof.get(dom)
        # the blending kernels are bound by instruction issue / the FP32 pipe (no dense contraction, no tensor-core
        # roof; their DRAM traffic is a third of the algorithmic bytes), the per-Gaussian stages by HBM.  The dominant kernel's
        # roofline is therefore the FP32 one; the HBM figure the contract asks for is kept in `hbm`.
        roofline = {"kernel": dom, "traffic": traffic,
                    "hbm": {"bound": "hbm", "achieved": stages[dom]["GBps"], "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                            "frac": stages[dom]["frac_hbm"]}}
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        mhz = (clocks or {}).get("sm_mhz") or 1965
        fp32_peak = sms * 128 * 2 * mhz * 1e6
        if isinstance(pairs, tuple):
            pairs_f, pairs_b, lanes_eval = pairs
            comp = {"fp32_peak_tflops": round(fp32_peak / 1e12, 1), "pairs_f": pairs_f, "pairs_b": pairs_b, "lanes_evaluated_fwd": lanes_eval,
                    "model": "16 flop per (pixel, entry) pair up to the last contributor (fwd), 60 per blended pair (bwd)"}
            for name, npairs, fl in (("blend_forward", pairs_f, 16), ("blend_backward", pairs_b, 60)):
                if name in stages:
                    tf = npairs * fl / (stages[name]["ms"] * 1e-3)
                    comp[name] = {"tflops": round(tf / 1e12, 2), "frac_fp32": round(tf / fp32_peak, 4)}
            comp["issue_active_pct"] = prof.get("issue_active_pct")
            roofline["compute"] = comp
            if dom in comp:
                roofline.update({"bound": "fp32 pipe / instruction issue (ncu: profiles/, issue_active_pct)", "achieved": comp[dom]["tflops"],
                                 "peak": comp["fp32_peak_tflops"], "peak_source": "SMs x 128 lanes x 2 x measured SM clock", "unit": "TFLOP/s",
                                 "frac": comp[dom]["frac_fp32"]})
            # the whole view against SURVEY 8(d)'s bound: t_min = sum(bytes) / HBM + 16 pairs_f / FP32 + 60 pairs_b / FP32
            t_hbm = sum(sb.values()) / (peak * 1e9) * 1e3
            t_f = pairs_f * 16 / fp32_peak * 1e3
            t_b = pairs_b * 60 / fp32_peak * 1e3
            t_meas = ms_total / args.steps / vps
            roofline["view"] = {"t_min_ms": round(t_hbm + t_f + t_b, 4), "t_hbm_ms": round(t_hbm, 4), "t_fp32_fwd_ms": round(t_f, 4),
                                "t_fp32_bwd_ms": round(t_b, 4), "t_measured_ms": round(t_meas, 4),
                                "achieved": round((t_hbm + t_f + t_b) / t_meas, 4),
                                "note": "SURVEY 8(d): roofline.achieved of the whole view = t_min / t_measured"}
        else:
            roofline.update(roofline["hbm"])
            roofline["compute"] = {"unavailable": pairs}
        cpu_baseline = None
        if world == 1 and not args.no_cpu_baseline:
            cpu_baseline = cpu_baseline_object(spec, P)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_total / args.steps, "ms_per_view": ms_total / args.steps / vps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "P": P, "W": spec.W, "H": spec.H, "sh_degree": spec.D, "views_per_step": vps * world,
                           "views_per_gpu_per_step": vps, "R_per_view": R, "visible_per_view": int(Pv),
                           "parallelism": "view-sharded dp%d, NCCL all-reduce of %d MB per step%s" % (
                               world, flat_nbytes >> 20, (", issued in %d ranges behind the last view's per-Gaussian backward" % overlap) if overlap else ""),
                           "grad_sum": ("deferred: blending backward per view, one per-Gaussian backward over the step's views "
                                        "(set_grad_mode('deferred') + flush_deferred)" if deferred else
                                        "in the backward kernels (set_grad_mode)" if fused else "autograd .grad +="),
                           "api": "GaussianRasterizerBatched: one call per rank and step" if args.batched else
                                  ("GaussianRasterizer: one call per view; the forward of view i+1 is issued before the backward of view i, its "
                                   "front (preprocess, depth order, tile split) on a high-priority stream (set_front_stream), the views alternating over "
                                   + str(n_fronts) + " such stream(s)" if piped else
                                   "GaussianRasterizer: one call per view"),
                           "c5": "config C5 (6M Gaussians, %d views per step) measured by the same step at this N: see the `c5` object" % (vps * world),
                           "l2": "inputs (%.1f GB of parameters + state per view) exceed the 126 MB L2; no flush needed" %
                                 ((flat_nbytes + 48 * P + 20 * R) / 1e9)},
                "roofline": roofline, "stages": stages, "stages_pipelined_ms": stages_piped, "cpu_baseline": cpu_baseline, "e2e": e2e, "allreduce": allreduce, "c5": c5,
                "reference_cuda": ref_extra, "plain_api": plain_extra, "grad_sum_in_kernels": kernels_extra, "dist2": dist2_extra,
                "batched": batched_extra, "one_stream": one_stream, "two_streams": streams_extra, "forward_only": forward_extra, "train_step": train_extra,
                "gpu_launches": launches, "device_allocs_in_timed_region": int(dev_allocs),
                "clocks": clocks}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def probe_R(C, params, cam, bg, D):
    """num_rendered of one view through the raw binding (untimed)."""
    import torch
    e = torch.Tensor([])
    with torch.no_grad():
        out = C.rasterize_gaussians(bg, params["means3D"], e, params["opacities"], params["scales"], params["rotations"], 1.0, e,
                                    cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, params["shs"], D,
                                    cam.campos, False, False)
    return int(out[0])


def probe_pairs(C, params, cam, bg, D):
    """(pairs_f, pairs_b, lanes evaluated) of one view, untimed: SURVEY 8(d)'s units of the blending kernels' FP32 roofline.
    pairs_f = sum over pixels of the list entries up to the pixel's last contributor (from n_contrib);
    pairs_b = (pixel, entry) pairs that were actually blended, counted by the statistics kernel of the TEST-ONLY arms
    library (gsr_b200.load_arms(); loaded for this untimed probe only, after the headline legs; loading it registers hooks
    and changes nothing on the default path)."""
    import ctypes
    import torch
    import gsr_b200
    e = torch.Tensor([])
    arms = gsr_b200.load_arms()
    prev = C.set_option("blend_stats", 1)
    try:
        arms.gsr_debug_blend_stats(None, 1)
        with torch.no_grad():
            out = C.rasterize_gaussians(bg, params["means3D"], e, params["opacities"], params["scales"], params["rotations"], 1.0, e,
                                        cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, params["shs"], D,
                                        cam.campos, False, False)
        torch.cuda.synchronize()
    finally:
        C.set_option("blend_stats", prev)
    st = (ctypes.c_ulonglong * 16)()
    arms.gsr_debug_blend_stats(st, 1)
    R, _color, _radii, geom, binning, img = out
    ncontrib = C.export_binning(geom, binning, img, int(params["means3D"].shape[0]), int(R), cam.W, cam.H)[3]
    return int(ncontrib.to(torch.int64).sum()), int(st[5]), int(st[2])


if __name__ == "__main__":
    main()
