#!/bin/bash
# r1d: 16-row chunks for k > 32 (build-time variant) against the shipped 32-row chunks
set -x
mkdir -p gpurun_out
for v in 32 16; do
echo "CHUNK_ROWS_LARGE=$v"
if [ $v = 16 ]; then cp modular-rag-system_b200/csrc/libragb200_c16.so modular-rag-system_b200/csrc/libragb200.so; fi
timeout 100 python -m scripts.perf_case 1000000 384 1024 100 fp16 20
timeout 200 python -m scripts.perf_case 12500000 384 4096 100 bf16 5
timeout 200 python -m scripts.perf_case 10000000 768 1024 100 fp16 5
done
(timeout 300 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q 2>&1 | tail -3)
