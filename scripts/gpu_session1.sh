#!/bin/bash
# one GPU session: parity tests, bench (both arms), ncu launch list, ncu full capture of the top kernel
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 3 --warmup 3 --no-extras --no-cpu > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_topk_pair -s 2 -c 1 -f -o gpurun_out/pair_10Mx768_b1024 \
  python -m scripts.prof_case 10000000 768 1024 4 > gpurun_out/ncu_full.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_topk_pair -s 2 -c 1 -f -o gpurun_out/pair_1Mx384_b1024 \
  python -m scripts.prof_case 1000000 384 1024 4 > gpurun_out/ncu_full2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:scan_topk -s 2 -c 1 -f -o gpurun_out/scan_10Mx768_b1 \
  python -m scripts.prof_case 10000000 768 1 4 > gpurun_out/ncu_full3.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json | head -c 3000
