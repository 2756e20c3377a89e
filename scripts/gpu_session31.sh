#!/bin/bash
# r1d: BASELINE config 5 at 8 GPUs with the final build
N=8
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29524 bench.py --gpus $N --steps 5 --warmup 3 --workload 100Mx384 --k 100 --batch 4096 > gpurun_out/r1d_bench_n${N}_100Mx384_top100_b4096.json 2> gpurun_out/r1d_bench_n$N.err; tail -1 gpurun_out/r1d_bench_n${N}_100Mx384_top100_b4096.json | cut -c1-300
