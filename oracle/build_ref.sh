#!/usr/bin/env bash
# Builds the UNMODIFIED reference CUDA extensions (diff-gaussian-rasterization and
# simple-knn) for sm_100 straight from the sources where they lie under
# /root/reference, into oracle/_ref/ (git-ignored, but shipped to the GPU box).
#
# This is TEST/BENCH infrastructure only: the resulting modules (ref_dgr_C,
# ref_knn_C) are the ground truth for the bit-exact integer contract and the
# "reference on the same B200" timing arm.  Nothing in the product imports them.
# No reference source is copied or modified; two missing standard includes are
# supplied with --pre-include (rasterizer_impl.h never includes <cstdint>,
# simple_knn.cu uses FLT_MAX without <cfloat>).
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
if [ ! -d "$REF/submodules/diff-gaussian-rasterization" ]; then
  echo "[build_ref] $REF not present - keeping prebuilt files in $OUT (if any)"; exit 0
fi
mkdir -p "$OUT/obj"
PY=${PYTHON:-python}
TORCH_INC=$($PY - <<'PYEOF'
import sysconfig, os, torch
from torch.utils.cpp_extension import include_paths
print(" ".join("-I"+p for p in include_paths()+[sysconfig.get_paths()["include"]]))
PYEOF
)
TORCH_LIB=$($PY -c "import torch,os;print(os.path.join(os.path.dirname(torch.__file__),'lib'))")
EXT=$($PY -c "import sysconfig;print(sysconfig.get_config_var('EXT_SUFFIX'))")
DGR="$REF/submodules/diff-gaussian-rasterization"
SK="$REF/submodules/simple-knn"
ARCH="-gencode arch=compute_100,code=sm_100"
# torch's BuildExtension adds exactly these to nvcc (no fast-math): keep numerics identical.
NVF="-std=c++17 -O3 $ARCH --expt-relaxed-constexpr -D__CUDA_NO_HALF_OPERATORS__ -D__CUDA_NO_HALF_CONVERSIONS__ -D__CUDA_NO_BFLOAT16_CONVERSIONS__ -D__CUDA_NO_HALF2_OPERATORS__ -Xcompiler -fPIC -I$DGR/third_party/glm"
CXXF="-std=c++17 -O2 -fPIC -w -D_GLIBCXX_USE_CXX11_ABI=1 -DTORCH_API_INCLUDE_EXTENSION_H $TORCH_INC -I/usr/local/cuda/include"

build_if_stale() { # out src cmd...
  local out=$1 src=$2; shift 2
  if [ ! -f "$out" ] || [ "$src" -nt "$out" ]; then echo "[build_ref] $out"; "$@"; fi
}
build_if_stale "$OUT/obj/forward.o"  "$DGR/cuda_rasterizer/forward.cu"  nvcc $NVF -c "$DGR/cuda_rasterizer/forward.cu"  -o "$OUT/obj/forward.o" &
build_if_stale "$OUT/obj/backward.o" "$DGR/cuda_rasterizer/backward.cu" nvcc $NVF -c "$DGR/cuda_rasterizer/backward.cu" -o "$OUT/obj/backward.o" &
build_if_stale "$OUT/obj/rasterizer_impl.o" "$DGR/cuda_rasterizer/rasterizer_impl.cu" nvcc $NVF --pre-include cstdint -c "$DGR/cuda_rasterizer/rasterizer_impl.cu" -o "$OUT/obj/rasterizer_impl.o" &
build_if_stale "$OUT/obj/simple_knn.o" "$SK/simple_knn.cu" nvcc $NVF --pre-include cfloat -c "$SK/simple_knn.cu" -o "$OUT/obj/simple_knn.o" &
# the two binding TUs contain no device code: compile them as plain C++ (minutes faster than nvcc)
build_if_stale "$OUT/obj/rasterize_points.o" "$DGR/rasterize_points.cu" g++ $CXXF -DTORCH_EXTENSION_NAME=ref_dgr_C -x c++ -c "$DGR/rasterize_points.cu" -o "$OUT/obj/rasterize_points.o" &
build_if_stale "$OUT/obj/dgr_ext.o" "$DGR/ext.cpp" g++ $CXXF -DTORCH_EXTENSION_NAME=ref_dgr_C -c "$DGR/ext.cpp" -o "$OUT/obj/dgr_ext.o" &
build_if_stale "$OUT/obj/spatial.o" "$SK/spatial.cu" g++ $CXXF -include cuda_runtime.h -DTORCH_EXTENSION_NAME=ref_knn_C -x c++ -c "$SK/spatial.cu" -o "$OUT/obj/spatial.o" &
build_if_stale "$OUT/obj/knn_ext.o" "$SK/ext.cpp" g++ $CXXF -DTORCH_EXTENSION_NAME=ref_knn_C -c "$SK/ext.cpp" -o "$OUT/obj/knn_ext.o" &
wait
LINK="-shared -L$TORCH_LIB -lc10 -ltorch_cpu -ltorch -ltorch_python -lc10_cuda -ltorch_cuda -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,$TORCH_LIB"
g++ "$OUT/obj/forward.o" "$OUT/obj/backward.o" "$OUT/obj/rasterizer_impl.o" "$OUT/obj/rasterize_points.o" "$OUT/obj/dgr_ext.o" $LINK -o "$OUT/ref_dgr_C$EXT"
g++ "$OUT/obj/simple_knn.o" "$OUT/obj/spatial.o" "$OUT/obj/knn_ext.o" $LINK -o "$OUT/ref_knn_C$EXT"
ls -la "$OUT"
