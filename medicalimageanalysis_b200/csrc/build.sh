#!/bin/bash
# Builds libmia_b200.so for sm_100a, in-tree (the .so travels to the GPU box with the snapshot).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="${MIA_OUT:-$HERE/../libmia_b200.so}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
    -Xcompiler -fPIC -Xcompiler -Wall -shared ${MIA_NVCC_EXTRA:-} \
    -o "$OUT" "$HERE/mia_b200.cu" -lcudart
echo "built $OUT"
