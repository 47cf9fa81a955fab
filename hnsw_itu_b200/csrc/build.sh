#!/bin/sh
# Builds libhnswb200.so in-tree for sm_100a (B200).  nvcc cross-compiles without a GPU.
set -e
here="$(cd "$(dirname "$0")" && pwd)"
out="${1:-$here/../libhnswb200.so}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -ccbin /usr/bin/g++ -Xcompiler -fPIC,-Wall,-Wno-unused-function -shared -cudart shared \
  -Xlinker -rpath,/usr/local/cuda/lib64 ${HB_NVCC_EXTRA} \
  -o "$out" "$here/api.cu" "$here/index.cu" "$here/search.cu" "$here/bruteforce.cu" "$here/synth.cu" "$here/builder.cu" "$here/idx_format.cpp"
echo "built $out"
