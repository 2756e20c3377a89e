#!/bin/bash
# top-10 (shared-memory candidate buffers): prune by selection (RAGB_PAIR_FULLSORT=2, default) against the bitonic top-32-of-64
# usage (GPU box): bash scripts/prune_ab.sh > gpurun_out/prune_ab.txt
for c in "1000000 384 1024 10 fp16" "1000000 768 1024 10 fp16" "10000000 384 1024 10 fp16" "10000000 768 1024 10 fp16" "1000000 384 128 10 fp16" "1000000 384 1024 32 fp16" "12500000 384 4096 100 bf16" "10000000 384 1024 64 fp16"; do
  set -- $c
  for w in 0 2; do
    echo -n "PRUNE=$w  "
    RAGB_PAIR_FULLSORT=$w python -m scripts.perf_case $1 $2 $3 $4 $5 10 2>&1 | tail -1
  done
done
