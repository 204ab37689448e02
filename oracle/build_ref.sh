#!/bin/bash
# Compiles the reference's AWQ kernels (csrc/quantization/awq/gemm_kernels.cu) and the two
# kernels on either side of the quantized linears (csrc/layernorm_kernels.cu,
# csrc/activation_kernels.cu), unmodified and from where they lie under /root/reference, for
# sm_100a into oracle/_ref/ref_awq.so (git-ignored, travels with gpurun).
# No reference source is copied into this repository.  ~6 minutes of nvcc.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
SRC="$REF/csrc/quantization/awq/gemm_kernels.cu"
OUT="$HERE/_ref"
[ -f "$SRC" ] || { echo "reference source not found: $SRC" >&2; exit 3; }
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
COMMON="-O2 -std=c++17 -DTORCH_EXTENSION_NAME=ref_awq -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=1 $INC -I$PY_INC -I$REF/csrc"
for f in "$SRC" "$REF/csrc/layernorm_kernels.cu" "$REF/csrc/activation_kernels.cu"; do
  nvcc -gencode arch=compute_100a,code=sm_100a $COMMON -Xcompiler -fPIC -c "$f" \
       -o "$OUT/$(basename "${f%.cu}").o" &
done
wait
for f in gemm_kernels layernorm_kernels activation_kernels; do [ -f "$OUT/$f.o" ] || exit 4; done
g++ $COMMON -fPIC -c "$HERE/ref_stub.cpp" -o "$OUT/ref_stub.o"
g++ -shared -o "$OUT/ref_awq$EXT_SUFFIX" "$OUT/gemm_kernels.o" "$OUT/layernorm_kernels.o" "$OUT/activation_kernels.o" "$OUT/ref_stub.o" \
    -L"$TORCH_LIB" -ltorch -ltorch_cpu -ltorch_python -lc10 -lc10_cuda -ltorch_cuda \
    -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,"$TORCH_LIB"
rm -f "$OUT"/*.o
echo "built $OUT/ref_awq$EXT_SUFFIX"
