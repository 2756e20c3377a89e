#!/bin/bash
for shape in "1000000 384" "100000 384" "10000000 768"; do
for b in 4 16 32 64 128; do
echo -n "TS   "; timeout 200 python -m scripts.perf_case $shape $b 10 fp16 10 2>&1 | tail -1
echo -n "PAIR "; RAGB_PAIR_SMALL=1 timeout 200 python -m scripts.perf_case $shape $b 10 fp16 10 2>&1 | tail -1
done; done
