#!/bin/sh
# Builds cli/hnsw-itu against the in-tree libhnswb200.so (rpath = the package directory).
set -e
here="$(cd "$(dirname "$0")" && pwd)"
lib="$here/../hnsw_itu_b200"
g++ -O2 -std=c++17 -Wall -o "$here/hnsw-itu" "$here/hnsw_itu.cpp" -L"$lib" -l:libhnswb200.so -Wl,-rpath,"$lib" -Wl,-rpath,/usr/local/cuda/lib64
echo "built $here/hnsw-itu"
