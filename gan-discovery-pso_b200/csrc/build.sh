#!/bin/bash
# Builds libpsob.so (sm_100a only) next to the Python host package.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="${HERE}/../psob/libpsob.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"${NVCC}" -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo \
  -Xcompiler -fPIC,-Wall,-Wno-unused-function -Xptxas -v --shared \
  -o "${OUT}" "${HERE}/swarm_kernels.cu" "${HERE}/pixel_gemm.cu" 2>&1
echo "built ${OUT}"
