#!/usr/bin/env bash
# TEST INFRASTRUCTURE — not part of the product.
#
# Compiles the UNMODIFIED reference rasterizer (yunjinli/TRASE, submodule
# diff-gaussian-rasterization) from the sources where they lie under
# /root/reference into oracle/_ref/.  Nothing is copied into the tracked tree:
# oracle/_ref/ is git-ignored (it still travels to the GPU box with gpurun).
#
# This does not run the reference's own build system (setup.py / CMake); it is
# four nvcc calls + one link, with the flags torch's CUDAExtension would pass
# (no -O flag, no fast-math, default -fmad=true) so that device code generation
# matches what a reference user gets.  `-include cstdint` is needed because
# cuda_rasterizer/rasterizer_impl.h:27 uses std::uintptr_t without <cstdint>
# (gcc 13).
#
# Result: oracle/_ref/diff_gaussian_rasterization/{__init__.py,_C.so}, oracle/_ref/trase_ref/utils/*.py
#   - _C.so exposes rasterize_gaussians / rasterize_gaussians_backward /
#     mark_visible (ext.cpp:15-19)
#   - __init__.py is the reference's own wrapper, installed (like
#     `pip install --target`) so the reference's public API can be driven on
#     the GPU box, where /root/reference does not exist.
set -euo pipefail
REF=${REF:-/root/reference/submodules/diff-gaussian-rasterization}
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/_ref/diff_gaussian_rasterization"
OBJ="$HERE/_ref/obj"
if [ ! -d "$REF" ]; then
  echo "build_ref: $REF not present (GPU box?) - keeping prebuilt oracle/_ref" >&2
  exit 0
fi
mkdir -p "$OUT" "$OBJ"
PY=${PYTHON:-python}
TORCH_INC=$($PY -c 'import os,torch;print(os.path.join(os.path.dirname(torch.__file__),"include"))' 2>/dev/null)
TORCH_LIB=$($PY -c 'import os,torch;print(os.path.join(os.path.dirname(torch.__file__),"lib"))' 2>/dev/null)
PY_INC=$($PY -c 'import sysconfig;print(sysconfig.get_paths()["include"])')
ARCH=${REF_ARCH:--gencode=arch=compute_100,code=sm_100}
FLAGS=(-D__CUDA_NO_HALF_OPERATORS__ -D__CUDA_NO_HALF_CONVERSIONS__
       -D__CUDA_NO_BFLOAT16_CONVERSIONS__ -D__CUDA_NO_HALF2_OPERATORS__
       --expt-relaxed-constexpr --compiler-options -fPIC -std=c++17
       -include cstdint -DTORCH_API_INCLUDE_EXTENSION_H
       -DTORCH_EXTENSION_NAME=_C -D_GLIBCXX_USE_CXX11_ABI=1
       -I"$REF/third_party/glm" -I"$REF"
       -isystem "$TORCH_INC" -isystem "$TORCH_INC/torch/csrc/api/include"
       -isystem "$PY_INC" -Xcompiler -fno-gnu-unique)
pids=()
for src in cuda_rasterizer/rasterizer_impl.cu cuda_rasterizer/forward.cu \
           cuda_rasterizer/backward.cu rasterize_points.cu; do
  o="$OBJ/$(basename "${src%.cu}").o"
  if [ ! -f "$o" ] || [ "$REF/$src" -nt "$o" ]; then
    nvcc "${FLAGS[@]}" $ARCH -c "$REF/$src" -o "$o" &
    pids+=($!)
  fi
done
if [ ! -f "$OBJ/ext.o" ] || [ "$REF/ext.cpp" -nt "$OBJ/ext.o" ]; then
  nvcc "${FLAGS[@]}" -x cu $ARCH -c "$REF/ext.cpp" -o "$OBJ/ext.o" &
  pids+=($!)
fi
for p in "${pids[@]}"; do wait "$p"; done
g++ -shared -o "$OUT/_C.so" "$OBJ"/*.o -L"$TORCH_LIB" -L/usr/local/cuda/lib64 \
    -lc10 -ltorch_cpu -ltorch -ltorch_python -lc10_cuda -ltorch_cuda -lcudart \
    -Wl,-rpath,"$TORCH_LIB"
install -m 0644 "$REF/diff_gaussian_rasterization/__init__.py" "$OUT/__init__.py"
# The reference's own Python for the pieces either side of the rasterizer that the reference arm of bench.py and the
# golden-vector scripts drive: DeformNetwork (utils/time_utils.py + rigid_utils.py) and the FEATURE-state / style
# losses (utils/loss_utils.py, utils/feature_utils.py).  Installed unmodified, as package `utils` under
# oracle/_ref/trase_ref/ (oracle/ref_loader.py:load_utils puts that directory on sys.path).
TOP="$(cd "$REF/../.." && pwd)"
mkdir -p "$HERE/_ref/trase_ref/utils"
: > "$HERE/_ref/trase_ref/utils/__init__.py"
for f in time_utils.py rigid_utils.py loss_utils.py feature_utils.py; do
  install -m 0644 "$TOP/utils/$f" "$HERE/_ref/trase_ref/utils/$f"
done
echo "build_ref: built $OUT/_C.so"
