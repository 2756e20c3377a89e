#!/bin/bash
for b in 16 128; do
RAGB_TS=1 timeout 200 python -m scripts.perf_case 10000000 768 $b 10 fp16 5 2>&1 | tail -1
RAGB_TS=1 RAGB_TS_KBS=2 timeout 200 python -m scripts.perf_case 10000000 768 $b 10 fp16 5 2>&1 | tail -1
done
RAGB_TS=1 timeout 200 python -m scripts.perf_case 1000000 384 128 10 fp16 20 2>&1 | tail -1
timeout 200 python -m scripts.perf_case 1000000 384 128 10 fp16 20 2>&1 | tail -1
