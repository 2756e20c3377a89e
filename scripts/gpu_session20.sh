#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for t in 8 4 16 32; do
echo "SYNC_TILES=$t"
RAGB_PAIR_SYNC_TILES=$t timeout 300 python -m scripts.perf_case 10000000 768 1024 10 fp16 20
RAGB_PAIR_SYNC_TILES=$t timeout 300 python -m scripts.perf_case 1000000 384 1024 10 fp16 50
done
RAGB_PAIR_SYNC_TILES=8 timeout 300 python -m scripts.perf_case 10000000 768 1024 10 fp16 20
env -u RAGB_PAIR_COOP timeout 300 ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum -k regex:gemm_topk_pair -s 2 -c 1 python -m scripts.prof_case 10000000 768 1024 4 2>&1 | grep -E "dram__bytes|duration"
