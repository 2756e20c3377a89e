#!/bin/bash
# r1d: sorted-prefix prune (k > 32) parity + A/B
set -x
mkdir -p gpurun_out
timeout 120 python -m scripts.perf_case 1000000 384 1024 100 fp16 20 || { echo "PRUNE FAILED/HUNG"; exit 1; }
(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > gpurun_out/r1d_pytest.log; cat gpurun_out/r1d_pytest.log
for q in 1 0; do
export RAGB_PAIR_FULLSORT=$q
echo "FULLSORT=$q"
timeout 300 python -m scripts.perf_case 12500000 384 4096 100 bf16 5
timeout 300 python -m scripts.perf_case 1000000 384 1024 100 fp16 20
timeout 300 python -m scripts.perf_case 1000000 384 1024 64 fp16 20
timeout 300 python -m scripts.perf_case 10000000 768 1024 100 fp16 5
done
