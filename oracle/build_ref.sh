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
