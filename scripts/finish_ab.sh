#!/bin/bash
# Finish kernel: chunks kept per query (RAGB_KSEL_SPARE: -1 = kpad, 2 = k + 2) x programmatic dependent launch (RAGB_FINISH_PDL)
# usage (GPU box): bash scripts/finish_ab.sh > gpurun_out/finish_ab.txt   — compare the ms/search column
for c in "1000000 384 1024 10 fp16" "10000000 768 1024 10 fp16" "1000000 384 1024 5 fp16" "1250000 768 1024 10 fp16"; do
  set -- $c
  for w in "-1 0" "-1 1" "2 0" "2 1"; do
    set -- $c $w
    echo -n "KSEL_SPARE=$6 PDL=$7  "
    RAGB_KSEL_SPARE=$6 RAGB_FINISH_PDL=$7 python -m scripts.perf_case $1 $2 $3 $4 $5 30 2>&1 | tail -1
  done
done
