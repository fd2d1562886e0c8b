#!/bin/bash
# Build the pybind11 / ATen host binding (csrc/fa_torch_ext.cpp) in-tree, next to libfa_b200.so:
#   vllm_flash_attn/_fa_torch.<abi>.so   — fwd / varlen_fwd / fwd_kvcache with the reference extension's signatures.
# Host-only C++ (no device code): g++ against the torch headers of the running interpreter, linked to libfa_b200.so.
set -e
cd "$(dirname "$0")"
PY=${PYTHON:-python}
read -r TORCH_DIR CXX11 EXT_SUFFIX PY_INC <<< "$($PY - <<'PYEOF'
import os, sysconfig, torch
print(os.path.dirname(torch.__file__), int(torch._C._GLIBCXX_USE_CXX11_ABI), sysconfig.get_config_var("EXT_SUFFIX"), sysconfig.get_paths()["include"])
PYEOF
)"
OUT=vllm_flash_attn/_fa_torch$EXT_SUFFIX
SRC=csrc/fa_torch_ext.cpp
if [ -f "$OUT" ] && [ "$OUT" -nt "$SRC" ] && [ "$OUT" -nt ../include/fa_b200.h ] && [ -z "$FA_FORCE" ]; then echo "up to date $OUT"; exit 0; fi
g++ -O2 -std=c++17 -fPIC -shared -DTORCH_EXTENSION_NAME=_fa_torch -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=$CXX11 \
    -I../include -I"$TORCH_DIR/include" -I"$TORCH_DIR/include/torch/csrc/api/include" -I"$PY_INC" -I/usr/local/cuda/include \
    $SRC -o "$OUT" \
    -L"$TORCH_DIR/lib" -ltorch -ltorch_cpu -ltorch_cuda -ltorch_python -lc10 -lc10_cuda -L/usr/local/cuda/lib64 -lcudart \
    -Lvllm_flash_attn -l:libfa_b200.so -Wl,-rpath,'$ORIGIN' -Wl,-rpath,"$TORCH_DIR/lib" -Wl,--no-as-needed
echo "built $OUT"
