#!/bin/bash
# Opportunistic prunes while the next accumulator is not there yet (RAGB_PAIR_SOFT=1, default) against forced prunes only
# usage (GPU box): bash scripts/soft_ab.sh > gpurun_out/soft_ab.txt
for c in "1000000 384 1024 10 fp16" "10000000 384 1024 10 fp16" "1000000 768 1024 10 fp16" "10000000 768 1024 10 fp16" "1000000 384 1024 100 fp16" "4000000 384 512 100 fp16" "12500000 384 4096 100 bf16" "1000000 384 1024 32 fp16"; do
  set -- $c
  for w in 0 1; do
    echo -n "SOFT=$w  "
    RAGB_PAIR_SOFT=$w python -m scripts.perf_case $1 $2 $3 $4 $5 20 2>&1 | tail -1
  done
done
