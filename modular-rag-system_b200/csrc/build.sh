#!/bin/bash
# Build libragb200.so in-tree for sm_100a.  Usage: csrc/build.sh [extra nvcc flags]
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
# RAGB_FAST_BUILD=1: split the optimiser over all cores (halves the build time; measured ~8% slower tcgen05 kernel:
# development only, never for a library that is benchmarked)
FAST=${RAGB_FAST_BUILD:+-split-compile 0}
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo \
  -Xcompiler -fPIC,-O3,-Wall -shared -cudart static $FAST \
  -o libragb200.so ragb200.cu "$@"
