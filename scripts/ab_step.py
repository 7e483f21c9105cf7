"""A/B of step-level variants inside ONE process (C3, 8 views per step, deferred + pipelined + prepared, the step of
bench.py's headline through gsr_b200.dp.deferred_views): the arms are run interleaved, several rounds, device-timed.
  arm = OPTION=value[;OPTION=value...]   (a name without '=' = defaults; options through gsr_debug_set_option, restored after the arm)
usage: python scripts/ab_step.py [--steps 10] [--rounds 3] [--arms=base,RADIX_DEPTH=1,FWD_MINB=4] [--out gpurun_out/ab_step.json]
Also reports every arm's gradients against the first arm's (equal up to the order of the atomic adds)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "structured-gaussian-splatting_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--arms", default="base,RADIX_DEPTH=1")
    ap.add_argument("--config", default="C3")
    ap.add_argument("--views-per-step", type=int, default=8)
    ap.add_argument("--out", default="gpurun_out/ab_step.json")
    args = ap.parse_args()
    import torch
    import diff_gaussian_rasterization as dgr
    from gsr_b200 import dp, synth
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    spec = synth.CONFIGS[args.config]
    scene = synth.spec_scene(spec, P=spec.P)
    params = dp.make_leaf_params(scene, dev)
    flat = dp.FlatGrads(params)
    V = 64
    cams = [synth.make_camera(k, V, spec.W, spec.H, spec.radius, spec.fovx_deg).to(dev) for k in range(V)]
    bg = synth.spec_background(spec, dev)
    go = synth.grad_image(spec.W, spec.H, device=dev)
    vps = args.views_per_step
    loss_fn = lambda color, k: (color * go).sum()

    def views(step):
        return [(step * vps + j) % V for j in range(vps)]

    fronts = [2]

    def step(i):
        dp.deferred_views(dgr, params, flat, cams, bg, spec.D, loss_fn, views(i), all_reduce=False, pipeline=True, prepare=True,
                          front_streams=fronts[0])

    def parse(a):
        return a, [(kv.split("=")[0], int(kv.split("=")[1])) for kv in a.split(";") if "=" in kv]       # (no '=': defaults)

    arms = [parse(a) for a in args.arms.split(",")]

    class applied:
        def __init__(self, opts):
            self.opts, self.prev = opts, []

        def __enter__(self):            # (FRONTS = n is not a library option: dp.deferred_views(front_streams=n))
            fronts[0] = dict(self.opts).get("FRONTS", 2)
            self.prev = [(k, dgr._C.set_option(k, v)) for k, v in self.opts if k != "FRONTS"]

        def __exit__(self, *exc):
            fronts[0] = 2
            for k, v in reversed(self.prev):
                dgr._C.set_option(k, v)

    res = {a[0]: [] for a in arms}
    ref_grad = None
    same = {}
    for name, opts in arms:                  # correctness: the same step under every arm
        with applied(opts):
            step(0)
            torch.cuda.synchronize()
        g = flat.flat.clone()
        if ref_grad is None:
            ref_grad = g
        same[name] = {"bit_identical": bool(torch.equal(g, ref_grad)),
                      "max_abs_diff": float((g - ref_grad).abs().max()), "ref_abs_max": float(ref_grad.abs().max())}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for r in range(args.rounds):
        for name, opts in arms:
            with applied(opts):
                for i in range(3):
                    step(i)
                torch.cuda.synchronize()
                e0.record()
                for i in range(args.steps):
                    step(3 + i)
                e1.record()
                torch.cuda.synchronize()
            res[name].append(e0.elapsed_time(e1) / (args.steps * vps))
    out = {"what": "ms per view, C3 step (deferred, pipelined, prepared), arms = option settings, %d rounds x %d steps"
                   % (args.rounds, args.steps),
           "ms_per_view": {k: [round(x, 4) for x in v] for k, v in res.items()},
           "best": {k: round(min(v), 4) for k, v in res.items()}, "same_gradients": same}
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
