#!/bin/sh
# Builds libhnswb200.so in-tree for sm_100a (B200): nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo (see Makefile).
set -e
here="$(cd "$(dirname "$0")" && pwd)"
out="${1:-$here/../libhnswb200.so}"
make -s -C "$here" -j"$(nproc)" OUT="$out"
