#!/bin/bash
# r1d: A/B switch test + stall reasons of the pair kernel's epilogue (top-100 at dim 384, top-10 at dim 384)
set -x
mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > gpurun_out/r1d_pytest.log; cat gpurun_out/r1d_pytest.log
for c in "top100 12500000 384 4096 2 100 bf16" "top10 1000000 384 1024 4 10 fp16"; do
set -- $c
name=r1d_stalls_$1
bash scripts/ncu_one.sh $name gemm_topk_pair 1 -- python -m scripts.prof_case $2 $3 $4 $5 $6 $7
( echo "# whole kernel"; python scripts/ncu_stall_reasons.py /tmp/$name.source.csv
  echo "# epilogue chunk path (gemm_pair.cuh lines 200-262 and 409-460)"; python scripts/ncu_stall_reasons.py /tmp/$name.source.csv 200 262; python scripts/ncu_stall_reasons.py /tmp/$name.source.csv 409 460
  echo "# prune / sort (lines 100-199, 263-300)"; python scripts/ncu_stall_reasons.py /tmp/$name.source.csv 100 199; python scripts/ncu_stall_reasons.py /tmp/$name.source.csv 263 300
) > gpurun_out/$name.txt 2>&1
done
rm -f gpurun_out/*.details.csv
cat gpurun_out/r1d_stalls_top100.txt
