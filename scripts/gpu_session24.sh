#!/bin/bash
# r1d evidence pass: tests, smoke, bench (both arms), launch list, ncu of the headline kernel and of norm_pack
set -x
mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > gpurun_out/r1d_pytest.log
(timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -3) > gpurun_out/r1d_smoke.log
timeout 600 python bench.py > gpurun_out/r1d_bench_n1.json 2> gpurun_out/r1d_bench_n1.err; echo "bench rc=$?"
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1d_bench_ref.json 2> gpurun_out/r1d_bench_ref.err
RAGB_PAIR_COOP=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 3 --warmup 3 --no-extras --no-cpu > gpurun_out/ncu_launch.log 2>&1
python scripts/launch_table.py gpurun_out/launches.csv > gpurun_out/r1d_launches_bench_10Mx768_b1024.txt
bash scripts/ncu_one.sh r1d_pair_10Mx768_b1024 gemm_topk_pair 2 -- python -m scripts.prof_case 10000000 768 1024 4
bash scripts/ncu_one.sh r1d_norm_pack_f32_f16_384 norm_pack 1 -- python scripts/norm_pack_perf.py
timeout 200 python scripts/norm_pack_perf.py
rm -f gpurun_out/*.details.csv gpurun_out/launches.csv
cat gpurun_out/r1d_pytest.log gpurun_out/r1d_smoke.log gpurun_out/r1d_launches_bench_10Mx768_b1024.txt
