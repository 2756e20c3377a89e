#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -m scripts.perf_case 10000000 768 1024 10 fp16 20
RAGB_PAIR_SYNC=0 timeout 300 python -m scripts.perf_case 10000000 768 1024 10 fp16 20
timeout 300 python -m scripts.perf_case 12500000 384 4096 100 bf16 5
RAGB_PAIR_SYNC=0 timeout 300 python -m scripts.perf_case 12500000 384 4096 100 bf16 5
timeout 300 ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum -k regex:gemm_topk_pair -s 2 -c 1 python -m scripts.prof_case 10000000 768 1024 4 2>&1 | grep -E "dram__bytes|duration"
