#!/bin/bash
set -x
mkdir -p gpurun_out
bash scripts/ncu_one.sh finish_1Mx384_b1024 finish_chunks 2 -- python -m scripts.prof_case 1000000 384 1024 4
cat gpurun_out/finish_1Mx384_b1024.ncu.txt; head -50 gpurun_out/finish_1Mx384_b1024.hot.txt
