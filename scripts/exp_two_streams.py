"""Gauge: forward+backward of independent C3 views alternating over two CUDA streams (two views in flight: the latency-bound
sort stages and the HBM-bound per-Gaussian kernels of one view next to the issue-bound blending kernels of the other),
against the same loop on one stream.  Separate parameter replicas per stream, so no gradient buffer is shared."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "structured-gaussian-splatting_b200"), ROOT]
import diff_gaussian_rasterization as dgr  # noqa: E402
from gsr_b200 import dp, synth  # noqa: E402


def main():
    spec = synth.CONFIGS["C3"]
    dev = torch.device("cuda")
    scene = synth.spec_scene(spec)
    cams = [synth.make_camera(k, 64, spec.W, spec.H, spec.radius, spec.fovx_deg).to(dev) for k in range(16)]
    bg = synth.spec_background(spec, dev)
    go = synth.grad_image(spec.W, spec.H, device=dev)
    NS = int(os.environ.get("NSTREAMS", 2))
    reps = [dp.make_leaf_params(scene, dev) for _ in range(NS)]
    flats = [dp.FlatGrads(p) for p in reps]
    streams = [torch.cuda.Stream() for _ in range(NS)]

    def one(params, k, mode):
        prev = dgr.set_grad_mode(mode)
        try:
            color, _r, _m = dp.render_view(dgr, params, cams[k % 16], bg, spec.D)
            color.backward(go)
        finally:
            dgr.set_grad_mode(prev)

    def run(n, multi):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if not multi:
            for k in range(n):
                one(reps[0], k, "overwrite" if k == 0 else "accumulate")
        else:
            cur = torch.cuda.current_stream()
            for s in streams:
                s.wait_stream(cur)
            for k in range(n):
                with torch.cuda.stream(streams[k % NS]):
                    one(reps[k % NS], k, "overwrite" if k < NS else "accumulate")
            for s in streams:
                cur.wait_stream(s)
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    res = {}
    for name, multi in (("one_stream_ms_per_view", False), ("multi_stream_ms_per_view", True), ("one_stream_again", False),
                        ("multi_stream_again", True)):
        run(4, multi)
        res[name] = run(32, multi)
    res["streams"] = NS
    print(json.dumps(res))


if __name__ == "__main__":
    main()
