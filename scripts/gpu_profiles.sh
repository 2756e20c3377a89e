#!/bin/bash
# Evidence pass for profiles/ (one gpurun call, one GPU): the ncu launch list of the bench command and one
# `ncu --set full` capture per hot kernel (summaries only come back; the reports stay on the box).
# Every command runs once without ncu first (its exit code decides), as the profiling recipe asks.
set -x
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-extras --no-cpu --no-parity --no-sustained"
$B > gpurun_out/plain_bench.log 2>&1 &&
RAGB_PAIR_COOP=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  $B > gpurun_out/ncu_launch.log 2>&1
python scripts/launch_table.py gpurun_out/launches.csv 12 > gpurun_out/launches_bench_10Mx768_b1024.txt
one() { name=$1; regex=$2; skip=$3; shift 3; "$@" > gpurun_out/plain_$name.log 2>&1 && bash scripts/ncu_one.sh $name $regex $skip -- "$@"; }
one pair_10Mx768_b1024 gemm_topk_pair 2 python -m scripts.prof_case 10000000 768 1024 4
one pair_1Mx384_b1024 gemm_topk_pair 2 python -m scripts.prof_case 1000000 384 1024 4
one pair_12p5Mx384_top100_b4096 gemm_topk_pair 4 python -m scripts.prof_case 12500000 384 4096 4 100 bf16
one solo_1Mx384_b128 gemm_topk_pair 2 python -m scripts.prof_case 1000000 384 128 4
one solo_10Mx768_b128 gemm_topk_pair 2 python -m scripts.prof_case 10000000 768 128 4
one finish_10Mx768_b1024 finish_chunks 2 python -m scripts.prof_case 10000000 768 1024 4
one finish_1Mx384_b1024 finish_chunks 2 python -m scripts.prof_case 1000000 384 1024 4
one scan_10Mx768_b1 scan_topk 2 python -m scripts.prof_case 10000000 768 1 4
one scan_10Mx384_b1 scan_topk 2 python -m scripts.prof_case 10000000 384 1 4
one scan_1Mx384_b1 scan_topk 2 python -m scripts.prof_case 1000000 384 1 4
one pool_fp16 pool_ 4 python scripts/ingest_perf.py
one norm_pack_f16_f16_384 norm_pack_sub 12 python scripts/norm_pack_perf.py
rm -f gpurun_out/*.details.csv
cat gpurun_out/launches_bench_10Mx768_b1024.txt
du -sh gpurun_out
