#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_ingest_gpu.py tests/test_rag_store_gpu.py -m gpu -x -q 2>&1 | tail -3
RAGB_NORM_PACK_GENERIC=1 timeout 200 python scripts/norm_pack_perf.py
timeout 200 python scripts/norm_pack_perf.py
timeout 300 python -m scripts.perf_case 12500000 384 4096 100 bf16 5
timeout 300 python -m scripts.perf_case 12500000 384 4096 10 bf16 5
timeout 300 python -m scripts.perf_case 1000000 384 1024 100 fp16 20
scripts/ncu_one.sh r1d_pair_12p5Mx384_top100_b4096 gemm_topk_pair 1 -- python -m scripts.prof_case 12500000 384 4096 2 100 bf16
python scripts/ncu_waits.py /tmp/r1d_pair_12p5Mx384_top100_b4096.ncu-rep > gpurun_out/r1d_pair_top100.waits.txt 2>&1 || true
