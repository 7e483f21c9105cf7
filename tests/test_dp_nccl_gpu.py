"""Multi-GPU correctness of the view-sharded data-parallel step with the REAL rasterizer: the flat gradient
buffer after dp_step on N ranks (NCCL all-reduce, plain and overlapped range by range, plain and NCCL-registered buffer)
equals the single-GPU gradient over the same views.  Needs >= 2 GPUs (skipped otherwise; world size 2)."""
import os
import sys

import numpy as np
import pytest
import torch

import util
from gsr_b200 import synth

pytestmark = pytest.mark.gpu

P, D, W, H, V = 20000, 3, 320, 200, 6


def _worker(rank, world, port, q):
    try:
        import torch.distributed as dist
        import diff_gaussian_rasterization as dgr
        from gsr_b200 import dp
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dp.init_nccl(dev, init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)      # (NCCL on a high-priority stream)
        sc, _, bg = util.small_scene(P, D, W, H, 111)
        bg = bg.to(dev)
        cams = [synth.make_camera(k, 9, W, H, 3.0, 50.0).to(dev) for k in range(V)]
        go = [synth.grad_image(W, H, seed=50 + k).to(dev) for k in range(V)]
        loss_fn = lambda color, k: (color * go[k]).sum()
        res = {}
        for name, kw in (("plain", dict()), ("overlap3", dict(overlap_chunks=3)), ("overlap5_registered", dict(overlap_chunks=5)),
                         ("deferred", dict()), ("deferred_overlap4_registered", dict(overlap_chunks=4)),
                         ("deferred_piped_prepared_overlap3", dict(overlap_chunks=3, pipeline=True, prepare=True))):
            params = dp.make_leaf_params(sc, dev)
            flat = dp.FlatGrads(params, nccl_registered=name.endswith("registered"))
            flat.flat.fill_(float("nan"))
            step = dp.deferred_views if name.startswith("deferred") else dp.dp_step
            total = step(dgr, params, flat, cams, bg, D, loss_fn, dp.shard_views(V, rank, world), **kw)
            torch.cuda.synchronize()
            res[name] = (flat.flat.clone().cpu().numpy(), float(total), bool(flat.registered))
            flat.release()
        ref = None
        if rank == 0:
            params = dp.make_leaf_params(sc, dev)
            flat = dp.FlatGrads(params)
            total = dp.backward_views(dgr, params, flat, cams, bg, D, loss_fn, range(V), fused=True)
            torch.cuda.synchronize()
            ref = (flat.flat.clone().cpu().numpy(), float(total))
        q.put((rank, res, ref, None))
        dist.barrier()
        dist.destroy_process_group()
    except Exception as ex:      # reported by the parent
        import traceback
        q.put((rank, None, None, traceback.format_exc()))


def test_dp_step_nccl_world2_equals_single_gpu():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs: p.start()
    out = sorted([q.get(timeout=600) for _ in procs], key=lambda x: x[0])
    for p in procs: p.join(120)
    for rank, res, ref, err in out:
        assert err is None, "rank %d failed:\n%s" % (rank, err)
    ref_flat, ref_total = out[0][2]
    assert np.isfinite(ref_flat).all()
    for name in ("plain", "overlap3", "overlap5_registered", "deferred", "deferred_overlap4_registered", "deferred_piped_prepared_overlap3"):
        f0, t0, _ = out[0][1][name]
        f1, t1, _ = out[1][1][name]
        assert np.array_equal(f0, f1), name + ": replicas must hold identical sums"
        assert util.rel_err(f0, ref_flat) <= (1e-4 if "piped" in name else 1e-5), name       # (atomics' order moves with concurrency)
        assert abs((t0 + t1) - ref_total) <= 1e-5 * max(1.0, abs(ref_total)), name
