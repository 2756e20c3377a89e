#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -4
for i in 1 2; do
for sync in 1 0; do
echo "RAGB_PAIR_SYNC=$sync"
RAGB_PAIR_SYNC=$sync timeout 300 python -m scripts.perf_case 10000000 768 1024 10 fp16 20
RAGB_PAIR_SYNC=$sync timeout 300 python -m scripts.perf_case 1000000 384 1024 10 fp16 50
RAGB_PAIR_SYNC=$sync timeout 300 python -m scripts.perf_case 12500000 384 4096 10 bf16 5
done
done
RAGB_PAIR_SYNC=1 bash scripts/ncu_one.sh pair_sync1 gemm_topk_pair 2 -- python -m scripts.prof_case 10000000 768 1024 4
grep -E "time_duration|dram__bytes_read|hit_rate|cycles_elapsed" gpurun_out/pair_sync1.ncu.txt
RAGB_PAIR_SYNC=0 bash scripts/ncu_one.sh pair_sync0 gemm_topk_pair 2 -- python -m scripts.prof_case 10000000 768 1024 4
grep -E "time_duration|dram__bytes_read|hit_rate|cycles_elapsed" gpurun_out/pair_sync0.ncu.txt
rm -f gpurun_out/*.details.csv
