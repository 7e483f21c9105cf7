"""In-tree build of the B200 rasterizer: CUDA core -> gsr_b200/libgsr_b200.so (plain C ABI, no torch),
torch binding -> gsr_b200/_binding*.so (host-only C++).  sm_100a only.

    python structured-gaussian-splatting_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU; objects are cached under build/ by source mtime.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
PKG = os.path.join(HERE, "gsr_b200")
OBJ = os.path.join(HERE, "build")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

CU_SOURCES = ["api.cu", "context.cu", "preprocess.cu", "preprocess_bwd.cu", "sort.cu", "blend.cu", "knn.cu", "loss.cu", "ssim.cu", "optim.cu", "io.cu", "activate.cu"]
# test-only comparison arms (plain blending kernels, statistics kernel, cp.async.bulk staging arm, CUB sort): a library of
# their own, loaded only by tests / benchmarks (gsr_b200.load_arms()); nothing of it is linked into libgsr_b200.so
ARM_SOURCES = [os.path.join("arms", "blend_arms.cu"), os.path.join("arms", "sort_cub.cu")]
HEADERS = ["gsr_common.cuh", "blend_common.cuh", "math3.cuh", os.path.join(INCLUDE, "gsr.h")]
# nvcc defaults on purpose (-fmad=true, IEEE div/sqrt, no fast-math): the bit-exact integer contract
# depends on the same fp32 rounding as the reference build (torch's BuildExtension adds no fast-math).
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo",
              "-Xcompiler", "-fPIC"]

LIB = os.path.join(PKG, "libgsr_b200.so")
ARMS_LIB = os.path.join(PKG, "libgsr_b200_arms.so")


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _mtime(p):
    return os.path.getmtime(p) if os.path.exists(p) else 0.0


def _stale(out, deps):
    t = _mtime(out)
    return t == 0.0 or any(_mtime(d) > t for d in deps)


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s" % (" ".join(cmd), r.stdout))
    if verbose and r.stdout.strip():
        print(r.stdout)


def binding_path():
    return os.path.join(PKG, "_binding" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_core(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    hdrs = [h if os.path.isabs(h) else os.path.join(CSRC, h) for h in HEADERS]
    jobs, objs = [], []
    for src in CU_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            jobs.append([_nvcc()] + NVCC_FLAGS + ["-c", s, "-o", o])
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(lambda c: _run(c, verbose), jobs))
    if force or jobs or _stale(LIB, objs):
        _run([_nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static"], verbose)
    return LIB


def build_arms(force=False, verbose=False):
    """libgsr_b200_arms.so: links against libgsr_b200.so (launch counter, device contexts, hook registration)."""
    os.makedirs(os.path.join(OBJ, "arms"), exist_ok=True)
    hdrs = [h if os.path.isabs(h) else os.path.join(CSRC, h) for h in HEADERS]
    jobs, objs = [], []
    for src in ARM_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            jobs.append([_nvcc()] + NVCC_FLAGS + ["-c", s, "-o", o])
    with ThreadPoolExecutor(max_workers=min(4, max(1, len(jobs)))) as ex:
        list(ex.map(lambda c: _run(c, verbose), jobs))
    if force or jobs or _stale(ARMS_LIB, objs + [LIB]):
        _run([_nvcc(), "-shared", "-o", ARMS_LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static",
              "-L", PKG, "-lgsr_b200", "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN"], verbose)
    return ARMS_LIB


def build_binding(force=False, verbose=False):
    out = binding_path()
    src = os.path.join(CSRC, "torch_binding.cpp")
    if not (force or _stale(out, [src, os.path.join(INCLUDE, "gsr.h")])):
        return out
    import torch
    from torch.utils.cpp_extension import include_paths
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    inc = []
    for p in include_paths() + [sysconfig.get_paths()["include"], "/usr/local/cuda/include"]:
        inc += ["-I", p]
    cmd = (["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-w", "-DTORCH_EXTENSION_NAME=_binding",
            "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)]
           + inc + [src, "-o", out, "-L", PKG, "-lgsr_b200", "-L", tlib, "-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_python",
                    "-lc10_cuda", "-ltorch_cuda", "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + tlib])
    _run(cmd, verbose)
    return out


def build(force=False, verbose=False):
    build_core(force, verbose)
    build_arms(force, verbose)
    build_binding(force, verbose)


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="--verbose" in sys.argv or "-v" in sys.argv)
    print("built", LIB, ARMS_LIB, "and", binding_path())
