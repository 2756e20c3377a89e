#!/bin/bash
# r1d: sorted-prefix flush parity + timing
set -x
mkdir -p gpurun_out
timeout 100 python -m scripts.perf_case 1000000 384 1024 100 fp16 20 || { echo "FLUSH FAILED/HUNG"; exit 1; }
(timeout 400 python -m pytest tests/test_gemm_gpu.py tests/test_sharded_gpu.py tests/test_search_gpu.py -m gpu -x -q 2>&1 | tail -3) > gpurun_out/r1d_pytest_flush.log; cat gpurun_out/r1d_pytest_flush.log
RAGB_PAIR_FULLSORT=1 timeout 100 python -m scripts.perf_case 1000000 384 1024 100 fp16 20
timeout 100 python -m scripts.perf_case 1000000 384 1024 64 fp16 20
timeout 200 python -m scripts.perf_case 12500000 384 4096 100 bf16 5
