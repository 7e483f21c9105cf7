"""`simple_knn._C` - distCUDA2 as exported by the reference (submodules/simple-knn/ext.cpp:15-17),
served by the B200 extension."""
from gsr_b200 import load_binding as _load

distCUDA2 = _load().distCUDA2
