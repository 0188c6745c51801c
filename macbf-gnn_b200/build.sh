#!/bin/bash
# Build libmacbf_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
OUT="$HERE/libmacbf_b200.so"
SRCS=("$HERE"/csrc/*.cu)
"$NVCC" -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a \
  -Xcompiler -fPIC -shared ${MACBF_NVCC_FLAGS:-} -o "$OUT" "${SRCS[@]}"
echo "built $OUT"
# measurement tool (tensor-core / FFMA peak probe); the binary travels to the GPU box with the snapshot
"$NVCC" -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -o "$HERE/tools/peak_probe" "$HERE/tools/peak_probe.cu"
echo "built $HERE/tools/peak_probe"
