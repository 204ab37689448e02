#!/bin/bash
# Builds integration/b200_shim.cpp (the pybind binding of INTEGRATION.md section A) against this image's
# torch and libb200_w4a16.so into integration/_build/ (git-ignored; travels with gpurun).  No GPU needed.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
OUT="$HERE/_build"
mkdir -p "$OUT"
PY="${PYTHON:-python}"
read -r TORCH_INC TORCH_LIB PY_INC EXT_SUFFIX <<<"$($PY - <<'PYEOF'
import sysconfig, os, torch
from torch.utils import cpp_extension as c
inc = " ".join("-I" + p for p in c.include_paths())
print(inc.replace(" ", ","), os.path.join(os.path.dirname(torch.__file__), "lib"),
      sysconfig.get_paths()["include"], sysconfig.get_config_var("EXT_SUFFIX"))
PYEOF
)"
INC="${TORCH_INC//,/ }"
LIBDIR="$ROOT/vllm-quick_b200/lib"
g++ -O2 -std=c++17 -fPIC -shared -DTORCH_EXTENSION_NAME=b200_shim -DTORCH_API_INCLUDE_EXTENSION_H \
    -D_GLIBCXX_USE_CXX11_ABI=1 $INC -I"$PY_INC" -I/usr/local/cuda/include -I"$ROOT/include" \
    "$HERE/b200_shim.cpp" -o "$OUT/b200_shim$EXT_SUFFIX" \
    -L"$LIBDIR" -lb200_w4a16 -L"$TORCH_LIB" -ltorch -ltorch_cpu -ltorch_python -lc10 -lc10_cuda -ltorch_cuda \
    -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,"$TORCH_LIB" -Wl,-rpath,'$ORIGIN/../../vllm-quick_b200/lib'
echo "built $OUT/b200_shim$EXT_SUFFIX"
