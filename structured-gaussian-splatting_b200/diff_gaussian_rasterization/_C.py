"""`diff_gaussian_rasterization._C` - same three entry points as the reference's pybind module
(submodules/diff-gaussian-rasterization/ext.cpp:15-19), served by the B200 extension."""
from gsr_b200 import load_binding as _load

_b = _load()
rasterize_gaussians = _b.rasterize_gaussians
rasterize_gaussians_backward = _b.rasterize_gaussians_backward
mark_visible = _b.mark_visible
export_binning = _b.export_binning          # test hook, not in the reference
launch_count = _b.launch_count
reset_launch_count = _b.reset_launch_count
