#!/bin/bash
# r1d multi-GPU pass: parity check under torchrun + the bench lines that changed this round
N=${1:-2}
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29520 scripts/check_multi_gpu.py 2>&1 | grep -v "^\*\|OMP_NUM" > gpurun_out/r1d_check_multi_gpu_n$N.log; tail -12 gpurun_out/r1d_check_multi_gpu_n$N.log
if [ "$N" = "8" ]; then
timeout 300 $TR --master-port 29521 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r1d_bench_n$N.json 2> gpurun_out/r1d_bench_n$N.err; tail -1 gpurun_out/r1d_bench_n$N.json | cut -c1-300
timeout 600 $TR --master-port 29524 bench.py --gpus $N --steps 5 --warmup 3 --workload 100Mx384 --k 100 --batch 4096 > gpurun_out/r1d_bench_n${N}_100Mx384_top100_b4096.json 2> gpurun_out/r1d_bench_n$N.err; tail -1 gpurun_out/r1d_bench_n${N}_100Mx384_top100_b4096.json | cut -c1-300
else
timeout 300 $TR --master-port 29521 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r1d_bench_n$N.json 2> gpurun_out/r1d_bench_n$N.err; tail -1 gpurun_out/r1d_bench_n$N.json | cut -c1-300
fi
tail -3 gpurun_out/r1d_bench_n$N.err
