#!/bin/bash
# r1d final evidence: bench (final build) + ncu of the top-100 kernel
set -x
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/r1d_bench_n1.json 2> gpurun_out/r1d_bench_n1.err; echo "bench rc=$?"
bash scripts/ncu_one.sh r1d_final_12p5Mx384_top100_b4096 gemm_topk_pair 1 -- python -m scripts.prof_case 12500000 384 4096 2 100 bf16
python scripts/ncu_waits.py /tmp/r1d_final_12p5Mx384_top100_b4096.ncu-rep > gpurun_out/r1d_final_top100.waits.txt 2>&1 || true
python scripts/ncu_stall_reasons.py /tmp/r1d_final_12p5Mx384_top100_b4096.source.csv > gpurun_out/r1d_final_top100.stalls.txt 2>&1 || true
rm -f gpurun_out/*.details.csv
cut -c1-300 gpurun_out/r1d_bench_n1.json
