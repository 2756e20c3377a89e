#!/bin/bash
# Scan kernel (K3), lanes per row A/B: RAGB_SCAN_WIDE=1 = widest group first (the r1 choice), 0 = fewest lanes (default).
# usage (GPU box): bash scripts/scan_lanes_ab.sh > gpurun_out/scan_lanes_ab.txt
for c in "1000000 384 fp16" "10000000 384 fp16" "1000000 768 fp16" "10000000 768 fp16" "4000000 1024 fp16" \
         "8000000 128 fp16" "8000000 256 fp16" "4000000 512 fp16" "4000000 384 fp32" "2000000 768 fp32" "10000000 384 bf16"; do
  set -- $c
  for w in 1 0; do
    echo -n "SCAN_WIDE=$w  "
    RAGB_SCAN_WIDE=$w python -m scripts.perf_case $1 $2 1 10 $3 30 2>&1 | tail -1
  done
done
