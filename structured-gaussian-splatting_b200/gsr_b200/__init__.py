"""Host-side helpers of the B200-native rasterizer hot path: loader for the C-ABI library and its
torch binding, synthetic scenes (synth), view-sharded data-parallel step (dp).

There is NO CPU or eager fallback: importing the binding fails loudly if the CUDA extension has not
been built (python structured-gaussian-splatting_b200/build.py, or __graft_entry__.build()).
"""
import os as _os

_HERE = _os.path.dirname(_os.path.abspath(__file__))
LIB_PATH = _os.path.join(_HERE, "libgsr_b200.so")


def load_binding():
    """Returns the torch extension module (gsr_b200._binding); raises ImportError if it is not built."""
    if not _os.path.exists(LIB_PATH):
        raise ImportError("gsr_b200: %s is missing - build the CUDA extension first "
                          "(python structured-gaussian-splatting_b200/build.py); there is no CPU fallback" % LIB_PATH)
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    from . import _binding
    return _binding


ARMS_PATH = _os.path.join(_HERE, "libgsr_b200_arms.so")
_ARMS = None


def load_arms():
    """Loads the TEST-ONLY comparison arms (plain blending kernels, statistics kernel, cp.async.bulk staging arm, CUB
    sort: csrc/arms/) and returns the ctypes handle.  Loading registers them with the core library; they are then
    selected through _binding.set_option("blend_arm" | "sort_cub" | "blend_stats", ...).  The product path never
    loads this library."""
    global _ARMS
    if _ARMS is None:
        import ctypes
        load_binding()                      # libgsr_b200.so must be mapped first
        if not _os.path.exists(ARMS_PATH):
            raise ImportError("gsr_b200: %s is missing (python structured-gaussian-splatting_b200/build.py)" % ARMS_PATH)
        _ARMS = ctypes.CDLL(ARMS_PATH, mode=ctypes.RTLD_GLOBAL)
    return _ARMS
