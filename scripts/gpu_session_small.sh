#!/bin/bash
# small-batch cases (4 <= B <= 128) at the headline corpus, the batch-2/3 scan, and a short bench line
mkdir -p gpurun_out
for b in 2 3 4 16 64 128 256; do
  timeout 300 python -m scripts.perf_case 10000000 768 $b 10 fp16 5 2>&1 | tail -1
done
for b in 4 64 128; do
  RAGB_PAIR_SMALL=1 timeout 300 python -m scripts.perf_case 10000000 768 $b 10 fp16 5 2>&1 | tail -1
done
timeout 600 python bench.py --steps 5 --warmup 3 --no-extras 2>/dev/null | grep "^{" | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['clocks'])"
