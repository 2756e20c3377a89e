#!/bin/bash
# full GPU suite + spot checks of the batch-size routing and the finish kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 200 python -m scripts.perf_case 1000000 384 128 10 fp16 20 2>&1 | tail -1
timeout 200 python -m scripts.perf_case 1000000 384 16 10 fp16 20 2>&1 | tail -1
timeout 200 python -m scripts.perf_case 1000000 384 64 100 fp16 20 2>&1 | tail -1
timeout 200 python -m scripts.perf_case 1000000 384 1024 10 fp16 20 2>&1 | tail -1
timeout 200 python -m scripts.perf_case 1000000 384 1024 100 fp16 20 2>&1 | tail -1
timeout 200 python -m scripts.perf_case 10000000 768 1024 10 fp16 5 2>&1 | tail -1
