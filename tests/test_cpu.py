"""CPU-only tests (no GPU): the oracle against the committed golden vectors (outputs of the unmodified
reference CUDA extension on a B200), the oracle's two independent backward derivations against each
other, the kNN oracle, the C-ABI library (loads, exports every symbol of include/gsr.h), host-side
API checks, and the view-sharded data-parallel plumbing over gloo with world size 2."""
import ctypes
import math
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

import util
from gsr_b200 import synth

ROOT = util.ROOT
GOLDEN = os.path.join(ROOT, "tests", "golden")


# ------------------------------------------------------------------------------ oracle vs golden (reference on B200)
def _oracle_vs_golden(name, sc, cam, bg, D):
    g = np.load(os.path.join(GOLDEN, "ref_%s.npz" % name))
    go = synth.grad_image(cam.W, cam.H, seed=6)
    st, og = util.run_oracle(sc, cam, bg, D, go)
    # integers: the CPU oracle does not fuse multiply-adds, so an ulp can flip a ceil(); count, don't hide
    mism = int((st.radii != g["radii"]).sum())
    assert mism <= max(1, st.P // 5000), "radii mismatches vs reference: %d" % mism
    if mism == 0:
        assert st.R == int(g["R"])
        assert np.array_equal(st.vals, g["vals"]), "sorted values"
        assert np.array_equal(st.keys, g["keys"]), "sorted keys"
        assert np.array_equal(st.ranges, g["ranges"]), "tile ranges"
        assert (st.n_contrib != g["n_contrib"]).sum() <= max(1, st.n_contrib.size // 10000)
    d = np.abs(st.color - g["color"]).max(axis=0)
    assert (d > 1e-4).sum() <= max(1, d.size // 10000), "image: %d pixels > 1e-4" % int((d > 1e-4).sum())
    pairs = [("means3D", "grad_means3D"), ("means2D", "grad_means2D"), ("opacities", "grad_opacity"), ("sh", "grad_sh"),
             ("scales", "grad_scales"), ("rotations", "grad_rotations")]
    for a, b in pairs:
        e = util.rel_err(og[a].reshape(-1), g[b].reshape(-1))
        assert e <= 1e-3, "%s: rel err %.3e vs reference" % (a, e)


def test_oracle_matches_reference_golden_c1():
    spec = synth.CONFIGS["C1"]
    _oracle_vs_golden("C1", synth.spec_scene(spec), synth.spec_camera(spec), synth.spec_background(spec), spec.D)


def test_oracle_matches_reference_golden_small_sh3():
    sc, cam, bg = util.small_scene(3000, 3, 200, 136, 41)
    _oracle_vs_golden("small3", sc, cam, bg, 3)


def test_oracle_backward_vs_autograd():
    """hand-derived backward (literal restatement of backward.cu) == torch autograd over the restated forward"""
    from oracle import torch_oracle
    sc, cam, bg = util.small_scene(300, 2, 48, 32, 3, ln_scale=math.log(0.06))
    go = synth.grad_image(48, 32, seed=9)
    st, og = util.run_oracle(sc, cam, bg, 2, go)
    color, ag = torch_oracle.render_autograd(st, go.numpy())
    assert np.abs(color - st.color).max() <= 2e-5
    vis = st.radii > 0
    for k, ref in (("means3D", ag["means3D"]), ("scales", ag["scales"]), ("rotations", ag["rotations"]), ("sh", ag["sh"])):
        assert util.rel_err(og[k][vis], ref[vis]) <= 2e-4, k
    assert util.rel_err(og["opacities"].reshape(-1)[vis], ag["opacities"][vis]) <= 2e-4
    assert util.rel_err(og["means2D"][vis][:, :2], ag["means2D"][vis]) <= 2e-4


def test_knn_oracle_morton_equals_brute_and_golden():
    import oracle
    g = torch.Generator().manual_seed(1)
    pts = (torch.rand(6000, 3, generator=g) * 2 - 1).numpy()
    pts[11] = pts[5]
    assert np.array_equal(oracle.dist2(pts, "morton"), oracle.dist2(pts, "brute"))
    gold = np.load(os.path.join(GOLDEN, "ref_knn20k.npz"))["dist2"]
    p20 = synth.make_scene(20000, 0, 1.0, seed=43)["means3D"].numpy()
    mine = oracle.dist2(p20, "morton")
    assert np.allclose(mine, gold, rtol=3e-7, atol=0)          # GPU fuses the distance FMAs, the CPU oracle does not


def test_sub_oracles_from_reference_python():
    """golden vectors produced by the reference's own torch restatements (utils/sh_utils.py eval_sh,
    scene/gaussian_model.py build_covariance_from_scaling_rotation), see tests/golden/make_sub_golden.py"""
    path = os.path.join(GOLDEN, "sub_sh_cov.npz")
    g = np.load(path)
    import oracle
    P = g["means"].shape[0]
    cam = synth.make_camera(0, 1, 64, 64, 3.0, 50.0)
    st = oracle.forward(g["means"], np.full((P, 1), 0.5, np.float32), shs=g["shs"], scales=g["scales"], rotations=g["rots"],
                        viewmatrix=cam.viewmatrix.numpy(), projmatrix=cam.projmatrix.numpy(), campos=g["campos"],
                        bg=np.zeros(3, np.float32), W=64, H=64, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, sh_degree=3,
                        scale_modifier=float(g["mod"]), blend=False)
    vis = st.radii > 0
    assert vis.sum() > P // 2
    assert np.abs(st.rgb[vis] - g["rgb"][vis]).max() <= 2e-6
    assert np.abs(st.cov3d[vis] - g["cov6"][vis]).max() <= 1e-6 * max(1.0, np.abs(g["cov6"]).max())


def test_loss_oracle_matches_reference_golden():
    """oracle/loss_oracle.py (numpy) against the reference's own l1_loss / ssim + torch autograd
    (utils/loss_utils.py:17-63, train.py:104-105), vectors from tests/golden/make_loss_golden.py"""
    from oracle import loss_oracle
    g = np.load(os.path.join(GOLDEN, "loss_l1_ssim.npz"))
    lam = float(g["lambda_dssim"])
    for k in "abc":
        o = loss_oracle.l1_ssim(g[k + "_image"], g[k + "_target"], lam)
        assert abs(o["l1"] - g[k + "_l1"]) <= 1e-6
        assert abs(o["ssim"] - g[k + "_ssim"]) <= 1e-6
        assert abs(o["loss"] - g[k + "_loss"]) <= 1e-6
        assert np.abs(o["grad"] - g[k + "_grad"]).max() <= 5e-6 * np.abs(g[k + "_grad"]).max()


def test_optim_oracle_matches_torch_adam_golden():
    """oracle/optim_oracle.py against torch.optim.Adam configured like scene/gaussian_model.py:164-175
    (vectors from tests/golden/make_optim_golden.py); densify_stats against the torch expressions of train.py:131-132"""
    from oracle import optim_oracle
    g = np.load(os.path.join(GOLDEN, "optim_adam.npz"))
    for n, lr in zip(g["names"], g["lrs"]):
        p = g["p0_" + n].copy(); m = np.zeros_like(p); v = np.zeros_like(p)
        for t in range(int(g["steps"])):
            optim_oracle.adam_step(p, g["g%d_%s" % (t, n)], m, v, t + 1, float(lr))
        assert np.abs(p - g["p_" + n]).max() <= 1e-7
        assert np.abs(m - g["m_" + n]).max() <= 1e-6 * np.abs(m).max()
        assert np.abs(v - g["v_" + n]).max() <= 1e-6 * np.abs(v).max()
    rng = np.random.default_rng(0)
    P = 200
    radii = rng.integers(-1, 30, P).astype(np.int32); radii[radii < 5] = 0
    dm = rng.standard_normal((P, 3)).astype(np.float32)
    mr = rng.uniform(0, 20, P).astype(np.float32); acc = rng.uniform(0, 1, P).astype(np.float32); den = np.ones(P, np.float32)
    tr, tm, ta, td = (torch.from_numpy(x.copy()) for x in (radii, mr, acc.reshape(P, 1), den.reshape(P, 1)))
    vis = tr > 0
    tm[vis] = torch.max(tm[vis], tr[vis].float())
    ta[vis] += torch.norm(torch.from_numpy(dm)[vis, :2], dim=-1, keepdim=True)
    td[vis] += 1
    optim_oracle.densify_stats(radii, dm, mr, acc, den)
    assert np.array_equal(mr, tm.numpy()) and np.array_equal(den, td.numpy().ravel())
    assert np.abs(acc - ta.numpy().ravel()).max() <= 1e-6


def _ply_model(P, deg, seed):
    rng = np.random.default_rng(seed)
    K = (deg + 1) ** 2 - 1
    f = lambda *s: rng.standard_normal(s).astype(np.float32)
    return dict(xyz=f(P, 3), features_dc=f(P, 1, 3), features_rest=f(P, K, 3), opacity=f(P, 1), scaling=f(P, 3), rotation=f(P, 4))


def test_ply_header_and_column_maps_match_oracle(tmp_path):
    """host side of gsr_b200.ply (header, column maps) against oracle/ply_oracle.py (scene/gaussian_model.py:187-266);
    the device gather out[r, j] = in[r, col[j]] is emulated with numpy indexing here and run for real in the gpu tests"""
    from gsr_b200 import ply
    from oracle import ply_oracle
    for deg, P in ((3, 41), (0, 5), (1, 0)):
        m = _ply_model(P, deg, 7 + deg)
        path = str(tmp_path / ("m%d.ply" % deg))
        ply_oracle.save_ply(path, **m)
        raw = open(path, "rb").read()
        n, props, off, fmt = ply.parse_header(raw[:1 << 16])
        K = (deg + 1) ** 2 - 1
        assert (n, fmt, len(props)) == (P, "binary_little_endian", 17 + 3 * K) and all(t == "f4" for _, t in props)
        assert raw[:off] == ply.format_header(P, ply.attribute_names(3 * K)) and len(raw) == off + 4 * P * len(props)
        payload, names = ply.read_payload(path)
        maps = ply.column_maps(names, deg)
        ref = ply_oracle.load_ply(path, deg)
        for k, cols in maps.items():
            assert np.array_equal(payload[:, cols].reshape(ref[k].shape), ref[k]) and np.array_equal(ref[k], m[k])
    with pytest.raises(ValueError):
        ply.column_maps(ply.attribute_names(45), 2)                 # the reference's assert on the f_rest count
    with pytest.raises(ValueError):
        ply.parse_header(b"plx\nend_header\n")
    # another exporter: reordered columns, a double and a uchar property, comments
    P = 9
    m = _ply_model(P, 1, 3)
    names = ply.attribute_names(9)
    order = list(reversed(names)) + ["extra"]
    dt = np.dtype([(k, "<f8" if k == "opacity" else ("u1" if k == "extra" else "<f4")) for k in order])
    el = np.zeros(P, dt)
    flat = np.concatenate((m["xyz"], np.zeros((P, 3), np.float32), m["features_dc"].transpose(0, 2, 1).reshape(P, -1),
                           m["features_rest"].transpose(0, 2, 1).reshape(P, -1), m["opacity"], m["scaling"], m["rotation"]), axis=1)
    for i, k in enumerate(names):
        el[k] = flat[:, i]
    path = str(tmp_path / "other.ply")
    with open(path, "wb") as f:
        f.write(b"ply\nformat binary_little_endian 1.0\ncomment made elsewhere\nelement vertex %d\n" % P)
        for k in order:
            f.write(("property %s %s\n" % ("double" if k == "opacity" else ("uchar" if k == "extra" else "float"), k)).encode())
        f.write(b"element face 0\nproperty list uchar int vertex_indices\nend_header\n")
        f.write(el.tobytes())
    payload, pn = ply.read_payload(path)
    maps = ply.column_maps(pn, 1)
    for k, cols in maps.items():
        assert np.array_equal(payload[:, cols].reshape(m[k].shape), m[k])


def test_activation_oracle_matches_torch_golden():
    """oracle/activation_oracle.py against torch.exp / F.normalize / torch.sigmoid + autograd (the reference's activation
    functions, scene/gaussian_model.py:33-40), vectors from tests/golden/make_activation_golden.py"""
    from oracle import activation_oracle as ao
    g = np.load(os.path.join(GOLDEN, "activations.npz"))
    y = ao.forward(g["scaling"], g["rotation"], g["opacity"])
    d = ao.backward(g["scaling"], g["rotation"], g["opacity"], g["g_scaling"], g["g_rotation"], g["g_opacity"])
    for a, k in zip(y, ("y_scaling", "y_rotation", "y_opacity")):
        assert np.abs(a - g[k]).max() <= 1e-6 * np.abs(g[k]).max()
    for a, k in zip(d, ("d_scaling", "d_rotation", "d_opacity")):
        assert np.abs(a - g[k]).max() <= 1e-6 * np.abs(g[k]).max()


def test_training_step_modules_import_without_gpu_and_refuse_cpu_tensors():
    """row f2-f4 host modules import on a CPU-only box (the binding is loaded lazily) and never fall back to the CPU"""
    from gsr_b200 import activations, losses, optim, ply, train_step   # noqa: F401
    for bad in (lambda: losses.l1_loss(torch.rand(4), torch.rand(4)),
                lambda: losses.l1_ssim_loss(torch.rand(3, 4, 4), torch.rand(3, 4, 4)),
                lambda: losses.ssim(torch.rand(3, 4, 4), torch.rand(3, 4, 4)),
                lambda: activations.activate(torch.zeros(2, 3), torch.zeros(2, 4), torch.zeros(2, 1)),
                lambda: ply.save_ply("/tmp/x.ply", *(torch.zeros(1, k) for k in (3, 3, 45, 1, 3, 4)))):
        with pytest.raises(RuntimeError):
            bad()
    p = torch.nn.Parameter(torch.ones(3))
    p.grad = torch.ones(3)
    with pytest.raises(RuntimeError):
        optim.FusedAdam([p], lr=0.1).step()
    assert [n for n, _ in train_step.PARAM_LR] == ["xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation"]


# ------------------------------------------------------------------------------ the C-ABI library
def test_biased_depth_keys_keep_the_order_and_leave_five_zero_bits():
    """The depth keys of radix_depth = 3 (csrc/preprocess.cu, GSR_DEPTH_KEY_BIAS in csrc/gsr_common.cuh): bits(depth) minus the
    bits of the near plane 0.2f.  Monotone over every depth the near cull lets through (z > 0.2 or NaN), never wrapping, and
    below 2^27 - five zero top bits, so the fourth radix pass is the identity - for every depth under 0.2 * 2^16."""
    bias = np.float32(0.2).view(np.uint32)
    assert int(bias) == 0x3E4CCCCD
    rng = np.random.default_rng(0)
    z = np.concatenate([np.nextafter(np.float32(0.2), np.float32(1.0), dtype=np.float32).reshape(1),
                        rng.uniform(0.2, 100.0, 200000).astype(np.float32), rng.uniform(100.0, 13107.0, 50000).astype(np.float32),
                        np.float32([13107.0, 13107.19, 13107.2, 13108.0, 1e6, 3e38, np.inf, np.nan])])
    z = z[~(z <= np.float32(0.2))]                        # the cull of preprocess: !(z <= 0.2) passes NaN like the reference
    bits = z.view(np.uint32)
    key = (bits - bias).astype(np.uint32)
    assert (bits >= bias).all()                           # no wrap-around
    o1 = np.argsort(bits, kind="stable"); o2 = np.argsort(key, kind="stable")
    assert np.array_equal(o1, o2)                         # the same stable order
    shallow = z < np.float32(0.2 * 65536.0)
    assert (key[shallow] < (1 << 27)).all() and (key[shallow] >= 1).all()
    assert (key[~shallow & ~np.isnan(z)] >= (1 << 27)).all()


def test_cabi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "gsr.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(gsr_[a-z0-9_]+)\s*\(", hdr))
    declared.discard("gsr_alloc_fn")
    assert {"gsr_forward", "gsr_backward", "gsr_mark_visible", "gsr_dist2", "gsr_export_binning"} <= declared
    lib = ctypes.CDLL(os.path.join(util.ROOT, "structured-gaussian-splatting_b200", "gsr_b200", "libgsr_b200.so"))
    for s in sorted(declared):
        assert hasattr(lib, s), "libgsr_b200.so does not export %s" % s
    lib.gsr_abi_version.restype = ctypes.c_int
    assert lib.gsr_abi_version() == 5
    lib.gsr_strerror.restype = ctypes.c_char_p
    assert b"precomputed" in lib.gsr_strerror(-4)
    lib.gsr_geom_bytes.restype = ctypes.c_size_t
    lib.gsr_binning_bytes.restype = ctypes.c_size_t
    assert lib.gsr_geom_bytes(1000) > 48 * 1000 and lib.gsr_binning_bytes(0, 5000, 64, 64) > 4 * 5000
    # argument validation happens before any CUDA call
    assert lib.gsr_forward(None, None, None, None, None, None) == -1
    assert lib.gsr_backward(None, None) == -1
    assert lib.gsr_forward_front(None, None, None, None, None, None, None) == -1
    assert lib.gsr_preprocess_views(None, 1, None, None) == -1
    assert lib.gsr_forward_prepared(None, None, None, None, None, None, None) == -6       # GSR_EBLOB: no prepared blob
    assert lib.gsr_backward_staged(None, 3, 0, -1, None) == -1
    assert lib.gsr_backward_views(None, 1, None, 0, -1, None) == -1
    # the product library carries no comparison arms (they live in the test-only libgsr_b200_arms.so) ...
    assert lib.gsr_debug_arms_loaded() == 0
    assert lib.gsr_debug_set_option(b"blend_arm", 1) == 0 and lib.gsr_debug_set_option(b"no_such_option", 1) == -1
    assert lib.gsr_debug_set_option(b"blend_arm", 0) == 0
    syms = subprocess.run(["nm", "-D", "--defined-only", os.path.join(util.ROOT, "structured-gaussian-splatting_b200", "gsr_b200",
                                                                      "libgsr_b200.so")], stdout=subprocess.PIPE, text=True).stdout
    assert "blend_fwd_v3_kernel" in syms and "blend_bwd_v6_kernel" in syms
    for retired in ("blend_stats_kernel", "blend_bwd_v4_kernel", "blend_bwd_v3_kernel", "cub_emit_kernel", "DeviceRadixSort"):
        assert retired not in syms, retired
    # ... and reads no environment variable on the hot path: getenv is referenced from context.cu's one-time reader only
    arms = ctypes.CDLL(os.path.join(util.ROOT, "structured-gaussian-splatting_b200", "gsr_b200", "libgsr_b200_arms.so"), mode=ctypes.RTLD_GLOBAL)
    assert hasattr(arms, "gsr_debug_blend_stats")


def test_python_api_surface_and_argument_checks():
    import diff_gaussian_rasterization as dgr
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")
    for name in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert callable(getattr(dgr._C, name))
    from simple_knn._C import distCUDA2
    assert callable(distCUDA2)
    cam = synth.make_camera(0, 1, 32, 32)
    rs = util.settings(dgr, cam, torch.zeros(3), 0, "cpu")
    r = dgr.GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=m)
    # no CPU fallback: CPU tensors are rejected loudly by the binding
    with pytest.raises(RuntimeError):
        r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=m, rotations=torch.zeros(4, 4))


# ------------------------------------------------------------------------------ data-parallel plumbing (gloo, world 2)
def _dp_worker(rank, world, port, q):
    import torch.distributed as dist
    from gsr_b200 import dp
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    sc = synth.make_scene(50, 1, seed=0)
    params = {k: sc[k].clone().requires_grad_(True) for k in dp.PARAM_ORDER}
    flat = dp.FlatGrads(params)
    views = dp.shard_views(6, rank, world)

    class FakeRast:         # a differentiable stand-in: the collective plumbing is what is under test
        class GaussianRasterizationSettings:
            def __init__(self, **kw): self.__dict__.update(kw)

        class GaussianRasterizer:
            def __init__(self, rs): self.rs = rs

            def __call__(self, means3D, means2D, opacities, shs, scales, rotations):
                s = (means3D.sum() + means2D.sum() + (opacities * self.rs.scale_modifier).sum() + shs.sum() * 2
                     + scales.sum() * 3 + rotations.sum() * 4)
                return s.expand(3, 2, 2), torch.ones(means3D.shape[0], dtype=torch.int32)

    cams = [synth.make_camera(k, 6, 32, 32) for k in range(6)]
    total = dp.dp_step(FakeRast, params, flat, cams, torch.zeros(3), 1, lambda color, k: color.sum() * (k + 1), views)
    q.put((rank, views, flat.flat.clone().numpy(), float(total)))
    dist.destroy_process_group()


def test_dp_gloo_world2():
    import torch.multiprocessing as mp
    from gsr_b200 import dp
    assert dp.shard_views(7, 0, 2) == [0, 2, 4, 6] and dp.shard_views(7, 1, 2) == [1, 3, 5]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs: p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda x: x[0])
    for p in procs: p.join(60)
    assert res[0][1] == [0, 2, 4] and res[1][1] == [1, 3, 5]
    assert np.array_equal(res[0][2], res[1][2])               # replicas hold identical summed gradients
    # sum over all 6 views of 12 * (k+1) * d(sum)/dparam:  means 1, scales 3, rotations 4, opacity 1, sh 2
    w = 12 * sum(range(1, 7))
    flat = res[0][2]
    P = 50
    assert np.allclose(flat[:3 * P], w * 1) and np.allclose(flat[3 * P:6 * P], w * 3) and np.allclose(flat[6 * P:10 * P], w * 4)
    assert np.allclose(flat[10 * P:11 * P], w * 1) and np.allclose(flat[11 * P:], w * 2)


# ------------------------------------------------------------------------------ the reference's callers, unchanged
REF_ROOT = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF_ROOT, "gaussian_renderer")), reason="reference checkout not present")
def test_reference_callers_import_and_call_our_packages_unchanged(monkeypatch):
    """north_star: "gaussian_renderer/__init__.py, train.py, train_lgm.py and render.py call it unchanged".  The reference's
    own gaussian_renderer (gaussian_renderer/__init__.py:14-100) is imported with OUR diff_gaussian_rasterization and simple_knn
    first on sys.path (plus a test-only `plyfile` stub: the trainer's un-vendored dependency, SURVEY section 7 hard part 7) and
    its render() is driven far enough to reach _C.rasterize_gaussians with the reference's 19 positional arguments.  No GPU
    here, so the module's hard-coded device="cuda" is redirected to the CPU and the binding call is recorded, not executed
    (the GPU parity tests cover what it computes)."""
    import types
    stub = types.ModuleType("plyfile")
    stub.PlyData = type("PlyData", (), {})
    stub.PlyElement = type("PlyElement", (), {})
    monkeypatch.setitem(sys.modules, "plyfile", stub)
    monkeypatch.syspath_prepend(REF_ROOT)                      # after ours: our packages stay first (conftest put them at the front)
    for name in [m for m in sys.modules if m.split(".")[0] in ("gaussian_renderer", "scene", "utils", "arguments")]:
        monkeypatch.delitem(sys.modules, name)
    sys.path.remove(REF_ROOT); sys.path.append(REF_ROOT)
    import diff_gaussian_rasterization as dgr
    import simple_knn._C as knn
    assert os.path.realpath(dgr.__file__).startswith(os.path.realpath(util.ROOT)) and hasattr(knn, "distCUDA2")
    import gaussian_renderer as gr                             # the reference's module, unmodified
    assert os.path.realpath(gr.__file__).startswith(REF_ROOT)
    assert gr.GaussianRasterizer is dgr.GaussianRasterizer and gr.GaussianRasterizationSettings is dgr.GaussianRasterizationSettings
    from scene.gaussian_model import GaussianModel
    import scene.gaussian_model as gm
    assert gm.distCUDA2 is knn.distCUDA2

    class TorchOnCpu:                                           # torch with device="cuda" requests served on the CPU
        def __getattr__(self, k):
            return getattr(torch, k)

        def zeros_like(self, *a, **kw):
            kw.pop("device", None)
            return torch.zeros_like(*a, **kw)

    monkeypatch.setattr(gr, "torch", TorchOnCpu())
    calls = []

    def fake_rasterize(*args):
        calls.append(args)
        P_ = args[1].shape[0]
        H_, W_ = args[12], args[13]
        z = torch.zeros(0, dtype=torch.uint8)
        return 0, torch.zeros(3, H_, W_), torch.ones(P_, dtype=torch.int32), z, z, z

    monkeypatch.setattr(dgr._C, "rasterize_gaussians", fake_rasterize)
    Pn = 12
    pc = GaussianModel(3)
    g = torch.Generator().manual_seed(0)
    pc._xyz = torch.randn(Pn, 3, generator=g).requires_grad_(True)
    pc._features_dc = torch.randn(Pn, 1, 3, generator=g).requires_grad_(True)
    pc._features_rest = torch.randn(Pn, 15, 3, generator=g).requires_grad_(True)
    pc._scaling = torch.randn(Pn, 3, generator=g).requires_grad_(True)
    pc._rotation = torch.randn(Pn, 4, generator=g).requires_grad_(True)
    pc._opacity = torch.randn(Pn, 1, generator=g).requires_grad_(True)
    pc.active_sh_degree = 2
    cam = synth.make_camera(0, 4, 64, 48)
    view = types.SimpleNamespace(FoVx=2 * math.atan(cam.tanfovx), FoVy=2 * math.atan(cam.tanfovy), image_height=48, image_width=64,
                                 world_view_transform=cam.viewmatrix, full_proj_transform=cam.projmatrix, camera_center=cam.campos)
    pipe = types.SimpleNamespace(convert_SHs_python=False, compute_cov3D_python=False, debug=False)
    out = gr.render(view, pc, pipe, torch.zeros(3))
    assert set(out) == {"render", "viewspace_points", "visibility_filter", "radii"}
    assert out["render"].shape == (3, 48, 64) and out["visibility_filter"].all() and out["radii"].shape == (Pn,)
    assert len(calls) == 1 and len(calls[0]) == 19
    (bg, means3D, colors, opac, scales, rots, mod, cov, vm, pm, tx, ty, Hh, Ww, sh, deg, campos, pref, dbg) = calls[0]
    assert means3D.shape == (Pn, 3) and colors.numel() == 0 and cov.numel() == 0 and sh.shape == (Pn, 16, 3)
    assert torch.allclose(scales, torch.exp(pc._scaling)) and torch.allclose(opac, torch.sigmoid(pc._opacity))
    assert torch.allclose(rots.norm(dim=1), torch.ones(Pn)) and (Hh, Ww, deg, mod, pref, dbg) == (48, 64, 2, 1.0, False, False)
    assert abs(tx - cam.tanfovx) < 1e-6 and torch.equal(vm, cam.viewmatrix) and torch.equal(pm, cam.projmatrix) and torch.equal(campos, cam.campos)
    # the python-side SH branch of render(): precomputed colours reach the binding as the reference sends them
    # (its compute_cov3D_python branch allocates on device="cuda" inside utils/general_utils.py:102 and cannot run here)
    pipe2 = types.SimpleNamespace(convert_SHs_python=True, compute_cov3D_python=False, debug=False)
    gr.render(view, pc, pipe2, torch.zeros(3))
    a2 = calls[1]
    assert a2[2].shape == (Pn, 3) and a2[14].numel() == 0 and a2[7].numel() == 0 and a2[4].shape == (Pn, 3) and a2[5].shape == (Pn, 4)
