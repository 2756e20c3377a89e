#!/bin/bash
# Where the warps of the pair kernel wait (build variant with cycle counters, -DRAGB_PAIR_STATS):
#   cd modular-rag-system_b200/csrc && make -j B=../../build_variants/stats OUT=../../build_variants/libragb200_stats.so EXTRA=-DRAGB_PAIR_STATS
# usage (GPU box): bash scripts/pair_stats.sh [cases...] > gpurun_out/pair_stats.txt
export RAGB200_LIB=$PWD/build_variants/libragb200_stats.so RAGB_PAIR_STATS_PRINT=1
if [ $# -eq 0 ]; then set -- "1000000 384 1024 10 fp16" "10000000 768 1024 10 fp16" "12500000 384 4096 100 bf16" "1000000 384 1024 100 fp16"; fi
for c in "$@"; do
  echo "== $c"
  python -m scripts.perf_case $c 3 2>&1 | tail -${TAILN:-14}
done
