"""`diff_gaussian_rasterization._C` - same three entry points as the reference's pybind module
(submodules/diff-gaussian-rasterization/ext.cpp:15-19: rasterize_gaussians,
rasterize_gaussians_backward, mark_visible), served by the B200 extension, plus its test / bench
hooks (export_binning, launch_count, timing_*)."""
from gsr_b200 import load_binding as _load

_b = _load()
for _name in dir(_b):
    if not _name.startswith("_"):
        globals()[_name] = getattr(_b, _name)
del _name
