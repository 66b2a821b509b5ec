#!/bin/bash
# Development helper: compiles ONE instantiation of the decoder kernel (seconds instead of the library's two minutes) and
# dumps its SASS, for instruction-level A/B work on a copy of the kernel header.
# usage: tools/sass_one.sh <directory holding ldpc_kernels_tmem.cuh and its includes> <output prefix> [BG SCALE FUSED MODE]
#        (defaults: 1 1 false 1 = BG1, generic 0.8 scaling, decode-only, Z = 384 instantiation)
set -e
DIR=$(realpath "$1"); OUT="$2"
BG=${3:-1}; SCALE=${4:-1}; FUSED=${5:-false}; MODE=${6:-1}
TMP=$(mktemp -d)
cat > "$TMP/inst.cu" <<EOT
#include <cstdio>
#include <cstring>
#define LDPC_TM_DEVICE_ONLY
#include "ldpc_kernels_tmem.cuh"
namespace ldpc_b200 {
template __global__ void ldpc_decode_tmem_kernel<$BG, $SCALE, $FUSED, $MODE>(const __grid_constant__ TmemParams);
}
EOT
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --expt-relaxed-constexpr -I"$DIR" -cubin -Xptxas -v \
  -o "$OUT.cubin" "$TMP/inst.cu" 2>&1 | grep -E "registers|spill|error" || true
cuobjdump -sass "$OUT.cubin" | grep -E "^\s+/\*[0-9a-f]{4,5}\*/" > "$OUT.sass"
wc -l "$OUT.sass"
rm -rf "$TMP"
