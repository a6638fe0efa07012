#!/bin/bash
# Builds sctagger_b200/libsctagger_b200.so for sm_100a (in-tree: the .so travels to the GPU box with the snapshot).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="${SCT_OUT:-$HERE/../libsctagger_b200.so}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
HOSTCXX="${SCT_HOSTCXX:-/usr/bin/g++}"
"$NVCC" -ccbin "$HOSTCXX" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
    -Xptxas -v -shared -Xcompiler -fPIC,-O3,-Wall,-pthread \
    -o "$OUT" "$HERE/sctagger_b200.cu" "$HERE/sct_align.cu" "$HERE/sct_wl.cu" "$HERE/sct_io.cpp" -lz -ldl "$@"
echo "built $OUT"
