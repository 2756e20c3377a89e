#!/bin/bash
# One query tile (4 <= B <= 128): solo form of the tcgen05 kernel (RAGB_SOLO=1, default) against the earlier routing
# (RAGB_SOLO=0: pair kernel beside a zero tile below N*D = 4e9, single-CTA row-granular kernel above).
# usage (GPU box): bash scripts/solo_ab.sh > gpurun_out/solo_ab.txt
for c in "1000000 384 128 10 fp16" "1000000 384 16 10 fp16" "1000000 768 128 10 fp16" "10000000 384 128 10 fp16" \
         "10000000 768 128 10 fp16" "10000000 768 8 10 fp16" "1000000 384 128 100 fp16" "10000000 768 128 100 bf16"; do
  set -- $c
  for w in 0 1; do
    echo -n "SOLO=$w  "
    RAGB_SOLO=$w python -m scripts.perf_case $1 $2 $3 $4 $5 30 2>&1 | tail -1
  done
done
