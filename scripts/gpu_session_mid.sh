#!/bin/bash
# mid-batch (B=128) at the headline corpus: ncu of the single-tile kernel and of the pair kernel with one padded tile
mkdir -p gpurun_out
bash scripts/ncu_one.sh mid_ts_b128 gemm_topk_ts_kernel 1 -- python -m scripts.prof_case 10000000 768 128 2
RAGB_PAIR_SMALL=1 bash scripts/ncu_one.sh mid_pair_b128 gemm_topk_pair_kernel 1 -- python -m scripts.prof_case 10000000 768 128 2
rm -f gpurun_out/*.details.csv
for st in 2 3 4; do RAGB_TS_STAGES=$st timeout 200 python -m scripts.perf_case 10000000 768 128 10 fp16 5 2>&1 | tail -1; done
RAGB_TS_NT=128 timeout 200 python -m scripts.perf_case 10000000 768 128 10 fp16 5 2>&1 | tail -1
head -24 gpurun_out/mid_ts_b128.ncu.txt
