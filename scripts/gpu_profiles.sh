#!/bin/bash
# Evidence pass for profiles/: bench JSON, ncu launch list of the same bench command, one `ncu --set full` capture
# per dominant kernel (summaries only come back; reports stay on the box).
set -x
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
RAGB_PAIR_COOP=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 3 --warmup 3 --no-extras --no-cpu > gpurun_out/ncu_launch.log 2>&1
python scripts/launch_table.py gpurun_out/launches.csv > gpurun_out/launches_bench_10Mx768_b1024.txt
bash scripts/ncu_one.sh pair_10Mx768_b1024 gemm_topk_pair 2 -- python -m scripts.prof_case 10000000 768 1024 4
bash scripts/ncu_one.sh pair_1Mx384_b1024 gemm_topk_pair 2 -- python -m scripts.prof_case 1000000 384 1024 4
bash scripts/ncu_one.sh finish_10Mx768_b1024 finish_chunks 2 -- python -m scripts.prof_case 10000000 768 1024 4
bash scripts/ncu_one.sh scan_10Mx768_b1 scan_topk 2 -- python -m scripts.prof_case 10000000 768 1 4
bash scripts/ncu_one.sh scan_1Mx384_b1 scan_topk 2 -- python -m scripts.prof_case 1000000 384 1 4
bash scripts/ncu_one.sh pool_fp16 pool_norm_pack 4 -- python scripts/ingest_perf.py
rm -f gpurun_out/*.details.csv
cat gpurun_out/launches_bench_10Mx768_b1024.txt
du -sh gpurun_out
