#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
{
timeout 300 python -m scripts.perf_case 12500000 384 4096 100 bf16 3
RAGB_PAIR_NOSPLIT=1 timeout 300 python -m scripts.perf_case 12500000 384 4096 100 bf16 3
timeout 300 python -m scripts.perf_case 12500000 384 4096 10 bf16 3
RAGB_PAIR_NOSPLIT=1 timeout 300 python -m scripts.perf_case 12500000 384 4096 10 bf16 3
timeout 300 python -m scripts.perf_case 10000000 768 1024 10 fp16 10
timeout 300 python -m scripts.perf_case 1000000 384 1024 10 fp16 20
timeout 300 python -m scripts.perf_case 1000000 384 1024 100 fp16 20
} 2>&1 | grep -v "^+" > gpurun_out/perf9.log
cat gpurun_out/perf9.log
