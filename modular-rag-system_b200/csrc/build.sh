#!/bin/bash
# Build libragb200.so in-tree for sm_100a.  Usage: csrc/build.sh [EXTRA="extra nvcc flags"]
set -e
cd "$(dirname "$0")"
make -j"$(nproc)" "$@"
