#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_1M_b1024.csv \
  python -m scripts.prof_case 1000000 384 1024 3 > gpurun_out/l1.log 2>&1
python scripts/launch_table.py gpurun_out/launches_1M_b1024.csv 3
timeout 300 python -m scripts.perf_case 1000000 384 1024 10 fp16 20
timeout 300 python -m scripts.perf_case 1000000 384 1024 100 fp16 20
timeout 300 python -m scripts.perf_case 10000000 768 1024 10 fp16 10
timeout 300 python -m scripts.perf_case 12500000 384 4096 100 bf16 3
