#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY (checker, never the product path).
# Compiles the reference's own CPU neighbor op from the sources WHERE THEY LIE under
# /root/reference (neighbors/neighbors.cpp:3-5 schema, neighbors/neighbors_cpu.cpp:19-98 impl)
# into oracle/_ref/torchmdnet_neighbors_ref.so.  No reference source is copied into the repo.
# The reference's own build system (setup.py / CUDAExtension) is NOT run.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
mkdir -p "$OUT"
if [ ! -d "$REF/torchmdnet/neighbors" ]; then
  echo "reference not present at $REF; keeping prebuilt files in $OUT" >&2
  exit 0
fi
# up to date?  (the CUDA half takes ~8 minutes under nvcc: rebuilt only when a reference source or this script changed)
fresh() {  # fresh <output>: 0 if it exists and is newer than every input
  [ -f "$1" ] || return 1
  for src in "$REF"/torchmdnet/neighbors/*.c* "$HERE/build_ref.sh"; do [ "$1" -nt "$src" ] || return 1; done
  return 0
}
WANT_CUDA=${BUILD_REF_CUDA:-}
if fresh "$OUT/torchmdnet_neighbors_ref.so" && [ -z "${FORCE:-}" ] && { [ -z "$WANT_CUDA" ] || fresh "$OUT/torchmdnet_neighbors_ref_cuda.so"; }; then
  echo "built $OUT/torchmdnet_neighbors_ref.so (up to date)"
  [ -n "$WANT_CUDA" ] && echo "built $OUT/torchmdnet_neighbors_ref_cuda.so (up to date)"
  exit 0
fi
PY=${PYTHON:-python}
read -r TORCH_INC TORCH_LIB PY_INC ABI <<<"$($PY - <<'PYEOF'
import torch, sysconfig, os
from torch.utils.cpp_extension import include_paths
print(";".join(include_paths()), os.path.join(os.path.dirname(torch.__file__), "lib"),
      sysconfig.get_paths()["include"], int(torch._C._GLIBCXX_USE_CXX11_ABI))
PYEOF
)"
INCS=""
IFS=';' read -ra P <<<"$TORCH_INC"
for p in "${P[@]}"; do INCS="$INCS -isystem $p"; done
g++ -O2 -std=c++17 -fPIC -shared -D_GLIBCXX_USE_CXX11_ABI=$ABI -DTORCH_EXTENSION_NAME=torchmdnet_neighbors_ref \
    $INCS -isystem "$PY_INC" \
    "$REF/torchmdnet/neighbors/neighbors.cpp" "$REF/torchmdnet/neighbors/neighbors_cpu.cpp" \
    -L"$TORCH_LIB" -ltorch -ltorch_cpu -lc10 -ltorch_python -Wl,-rpath,"$TORCH_LIB" \
    -o "$OUT/torchmdnet_neighbors_ref.so"
echo "built $OUT/torchmdnet_neighbors_ref.so"

# ---- GPU-side baseline (bench/test infrastructure only): the reference's own CUDA neighbour kernels
# (neighbors/neighbors_cuda.cu + neighbors_cuda_{brute,shared,cell}.cuh) recompiled for sm_100 from where they lie,
# into oracle/_ref/torchmdnet_neighbors_ref_cuda.so.  It registers the SAME op name as the product's shim, so it is only
# ever loaded in a process of its own (neighbor_sweep.py --impl reference).  nvcc cross-compiles without a GPU.
# Opt-in (BUILD_REF_CUDA=1 bash oracle/build_ref.sh): torch's headers cost nvcc about 8 minutes.
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
if [ -n "$WANT_CUDA" ] && [ -x "$NVCC" ]; then
  NINCS=""
  for p in "${P[@]}"; do NINCS="$NINCS -isystem $p"; done
  "$NVCC" -O3 -std=c++17 -shared -Xcompiler -fPIC -gencode arch=compute_100,code=sm_100 \
      -D_GLIBCXX_USE_CXX11_ABI=$ABI -DTORCH_EXTENSION_NAME=torchmdnet_neighbors_ref_cuda \
      -D__CUDA_NO_HALF_OPERATORS__ -D__CUDA_NO_HALF_CONVERSIONS__ -D__CUDA_NO_HALF2_OPERATORS__ --expt-relaxed-constexpr \
      $NINCS -isystem "$PY_INC" \
      "$REF/torchmdnet/neighbors/neighbors.cpp" "$REF/torchmdnet/neighbors/neighbors_cpu.cpp" \
      "$REF/torchmdnet/neighbors/neighbors_cuda.cu" \
      -L"$TORCH_LIB" -ltorch -ltorch_cpu -ltorch_cuda -lc10 -lc10_cuda -ltorch_python -Xlinker -rpath -Xlinker "$TORCH_LIB" \
      -o "$OUT/torchmdnet_neighbors_ref_cuda.so" 2> "$OUT/build_cuda.log" \
    && echo "built $OUT/torchmdnet_neighbors_ref_cuda.so" \
    || { echo "reference CUDA op did not build (see $OUT/build_cuda.log)" >&2; tail -5 "$OUT/build_cuda.log" >&2; }
fi
