#!/bin/bash
# full GPU suite + the 2/3-query cases at the headline corpus
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for b in 1 2 3; do
  timeout 300 python -m scripts.perf_case 10000000 768 $b 10 fp16 5 2>&1 | tail -1
done
