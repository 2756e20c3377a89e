#!/bin/bash
# scaling evidence: bench.py at N GPUs (headline, batch 1; at N=8 also the NCCL exchange, BASELINE config 5 and the parity check)
N=${1:-8}
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29521 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -1 gpurun_out/bench_n$N.json | cut -c1-300
timeout 300 $TR --master-port 29523 bench.py --gpus $N --steps 50 --warmup 3 --batch 1 > gpurun_out/bench_n${N}_b1.json 2>> gpurun_out/bench_n$N.err; tail -1 gpurun_out/bench_n${N}_b1.json | cut -c1-300
if [ "$N" = "8" ]; then
timeout 300 $TR --master-port 29520 scripts/check_multi_gpu.py 2>&1 | grep -v "^\*\|OMP_NUM" > gpurun_out/check_n$N.log; tail -30 gpurun_out/check_n$N.log
timeout 300 $TR --master-port 29522 bench.py --gpus $N --steps 20 --warmup 3 --exchange nccl > gpurun_out/bench_n${N}_nccl.json 2>> gpurun_out/bench_n$N.err; tail -1 gpurun_out/bench_n${N}_nccl.json | cut -c1-300
timeout 600 $TR --master-port 29524 bench.py --gpus $N --steps 5 --warmup 3 --workload 100Mx384 --k 100 --batch 4096 > gpurun_out/bench_n${N}_c5.json 2>> gpurun_out/bench_n$N.err; tail -1 gpurun_out/bench_n${N}_c5.json | cut -c1-300
fi
tail -3 gpurun_out/bench_n$N.err
