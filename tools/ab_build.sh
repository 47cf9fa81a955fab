#!/bin/bash
# Build search-kernel variants into variants/<tag>.so:  tools/ab_build.sh tag "-DHB_...=.." [tag flags ...]
set -e
cd "$(dirname "$0")/.."
mkdir -p variants
while [ $# -ge 2 ]; do
  tag="$1"; flags="$2"; shift 2
  make -s -C hnsw_itu_b200/csrc -j"$(nproc)" OUT="$PWD/variants/$tag.so" OBJDIR="$PWD/variants/obj_$tag" EXTRA="$flags" 2>&1 | grep -E "error|built" || true
done
