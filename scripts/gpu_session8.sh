#!/bin/bash
set -x
mkdir -p gpurun_out
bash scripts/ncu_one.sh pair_c5_k100 gemm_topk_pair 2 -- python -m scripts.prof_case 12500000 384 4096 4 100 bf16
cat gpurun_out/pair_c5_k100.ncu.txt; head -45 gpurun_out/pair_c5_k100.hot.txt | cut -c1-220
bash scripts/ncu_one.sh finish_c5_k100 finish_chunks 2 -- python -m scripts.prof_case 12500000 384 4096 4 100 bf16
cat gpurun_out/finish_c5_k100.ncu.txt; head -30 gpurun_out/finish_c5_k100.hot.txt | cut -c1-220
rm -f gpurun_out/*.details.csv
