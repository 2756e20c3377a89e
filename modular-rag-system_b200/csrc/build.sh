#!/bin/bash
# Build libragb200.so in-tree for sm_100a.  Usage: csrc/build.sh [extra nvcc flags]
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo \
  -Xcompiler -fPIC,-O3,-Wall -shared -cudart static \
  -o libragb200.so ragb200.cu "$@"
