#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
{
timeout 300 python -m scripts.perf_case 1000000 384 1024 10 fp16 20
timeout 300 python -m scripts.perf_case 1000000 384 1 10 fp16 50
RAGB_SCAN_2PASS=1 timeout 300 python -m scripts.perf_case 1000000 384 1 10 fp16 50
timeout 300 python -m scripts.perf_case 10000000 768 1024 10 fp16 5
timeout 300 python -m scripts.perf_case 10000000 768 1 10 fp16 20
timeout 300 python -m scripts.perf_case 12500000 384 4096 100 bf16 3
timeout 300 python -m scripts.perf_case 12500000 384 4096 10 bf16 3
timeout 300 python -m scripts.perf_case 1000000 384 1024 100 fp16 10
} > gpurun_out/perf3.log 2>&1
cat gpurun_out/perf3.log
bash scripts/ncu_one.sh pair_1Mx384_b1024 gemm_topk_pair 2 -- python -m scripts.prof_case 1000000 384 1024 4
cat gpurun_out/pair_1Mx384_b1024.ncu.txt
