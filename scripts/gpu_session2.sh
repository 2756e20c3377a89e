#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -m scripts.perf_case 12500000 384 4096 100 bf16 3 > gpurun_out/perf_c5.log 2>&1
timeout 300 python -m scripts.perf_case 12500000 384 4096 10 bf16 3 >> gpurun_out/perf_c5.log 2>&1
timeout 300 python -m scripts.perf_case 1000000 384 1024 100 fp16 10 >> gpurun_out/perf_c5.log 2>&1
cat gpurun_out/perf_c5.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 3 --warmup 3 --no-extras --no-cpu > gpurun_out/ncu_launch.log 2>&1
bash scripts/ncu_one.sh pair_10Mx768_b1024 gemm_topk_pair 2 -- python -m scripts.prof_case 10000000 768 1024 4
bash scripts/ncu_one.sh pair_1Mx384_b1024 gemm_topk_pair 2 -- python -m scripts.prof_case 1000000 384 1024 4
bash scripts/ncu_one.sh scan_10Mx768_b1 scan_topk 2 -- python -m scripts.prof_case 10000000 768 1 4
bash scripts/ncu_one.sh scan_1Mx384_b1 scan_topk 2 -- python -m scripts.prof_case 1000000 384 1 4
cat gpurun_out/pair_1Mx384_b1024.ncu.txt gpurun_out/pair_1Mx384_b1024.hot.txt | head -80
du -sh gpurun_out
