#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python scripts/ingest_perf.py > gpurun_out/ingest_perf.log 2>&1; cat gpurun_out/ingest_perf.log
bash scripts/ncu_one.sh pool_fp16 pool_norm_pack 4 -- python scripts/ingest_perf.py
cat gpurun_out/pool_fp16.ncu.txt; head -40 gpurun_out/pool_fp16.hot.txt
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
cat gpurun_out/bench.json
