#!/bin/bash
# usage: scripts/ncu_one.sh <name> <kernel-regex> <skip> -- <command...>
# One `ncu --set full` capture of one launch; the .ncu-rep stays on the box (too big for gpurun_out/),
# the raw-page summary, the per-section details and the hottest source lines come back as text.
name=$1; regex=$2; skip=$3; shift 4
rep=/tmp/$name
export RAGB_PAIR_COOP=0   # ncu kernel replay cannot measure cooperative launches; the kernel itself is unchanged
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 -f -o $rep "$@" > gpurun_out/$name.ncu.log 2>&1
python scripts/ncu_summary.py $rep.ncu-rep > gpurun_out/$name.ncu.txt 2>> gpurun_out/$name.ncu.log
ncu -i $rep.ncu-rep --page details --csv > gpurun_out/$name.details.csv 2>> gpurun_out/$name.ncu.log
ncu -i $rep.ncu-rep --page source --print-source cuda,sass --csv > /tmp/$name.source.csv 2>> gpurun_out/$name.ncu.log
python scripts/ncu_hot_lines.py /tmp/$name.source.csv > gpurun_out/$name.hot.txt 2>> gpurun_out/$name.ncu.log
ls -la $rep.ncu-rep >> gpurun_out/$name.ncu.log
