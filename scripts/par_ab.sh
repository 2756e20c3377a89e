#!/bin/bash
# k > 32 epilogue A/B: RAGB_PAIR_PAR (1 = tile-parity split with private candidate buffers, 0 = column split) x
# RAGB_PAIR_FULLSORT (2 = prune by selection, 0 = sorted-prefix bitonic prune).
# usage (GPU box): bash scripts/par_ab.sh > gpurun_out/par_ab.txt
for c in "12500000 384 4096 100 bf16" "1000000 384 1024 100 fp16" "10000000 768 1024 100 fp16" "10000000 384 1024 64 fp16" "4000000 384 512 100 fp16"; do
  set -- $c
  for w in "0 0" "0 2" "1 0" "1 2"; do
    set -- $c $w
    echo -n "PAR=$6 PRUNE=$7  "
    RAGB_PAIR_PAR=$6 RAGB_PAIR_FULLSORT=$7 python -m scripts.perf_case $1 $2 $3 $4 $5 10 2>&1 | tail -1
  done
done
