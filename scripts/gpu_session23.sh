#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 120 python -m scripts.perf_case 1000000 384 1024 100 fp16 20 || { echo "GTHR FAILED/HUNG"; exit 1; }
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for q in 0 1; do
export RAGB_PAIR_GTHR=$q
echo "GTHR=$q"
timeout 300 python -m scripts.perf_case 12500000 384 4096 100 bf16 5
timeout 300 python -m scripts.perf_case 12500000 384 4096 10 bf16 5
timeout 300 python -m scripts.perf_case 1000000 384 1024 100 fp16 20
timeout 300 python -m scripts.perf_case 1000000 384 1024 10 fp16 50
timeout 300 python -m scripts.perf_case 1000000 384 128 10 fp16 50
timeout 300 python -m scripts.perf_case 10000000 768 1024 10 fp16 10
timeout 300 python -m scripts.perf_case 10000000 768 1024 100 fp16 5
done
unset RAGB_PAIR_GTHR
timeout 200 python scripts/norm_pack_perf.py
scripts/ncu_one.sh r1d_pairg_12p5Mx384_top100_b4096 gemm_topk_pair 1 -- python -m scripts.prof_case 12500000 384 4096 2 100 bf16
python scripts/ncu_waits.py /tmp/r1d_pairg_12p5Mx384_top100_b4096.ncu-rep > gpurun_out/r1d_pairg_top100.waits.txt 2>&1 || true
