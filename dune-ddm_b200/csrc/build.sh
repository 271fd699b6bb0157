#!/bin/sh
# Builds libddm_b200.so in-tree for sm_100a (no other architecture is supported).
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
OUT="${DDM_OUT:-$HERE/../libddm_b200.so}"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a \
  -Xcompiler -fPIC -shared ${DDM_PTXAS_V:+-Xptxas -v} $DDM_NVCC_EXTRA \
  -o "$OUT" "$HERE/ddm_b200.cu" -lcusolver -lcublas -ldl
echo "built $OUT"
