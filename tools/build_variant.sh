#!/bin/bash
# Development helper for A/B runs: builds the product library with an alternative kernel header.
# usage: tools/build_variant.sh <kernel header to use as ldpc_kernels_tmem.cuh> <output .so>
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$(realpath "$2")
EXTRA="${@:3}"
TMP=$(mktemp -d)
mkdir -p "$TMP/pkg" "$TMP/include"
cp -r "$ROOT/srsran_project_23.5_b200/csrc" "$TMP/pkg/csrc"
cp "$ROOT/include/ldpc_cuda.h" "$TMP/include/"
cp "$1" "$TMP/pkg/csrc/ldpc_kernels_tmem.cuh"
cd "$TMP/pkg"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --expt-relaxed-constexpr $EXTRA -Xcompiler -fPIC -shared -o "$OUT" csrc/ldpc_api.cu csrc/ldpc_geometry.cpp csrc/ldpc_pool.cpp
rm -rf "$TMP"
