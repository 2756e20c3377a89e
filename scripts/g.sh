#!/bin/bash
# usage: scripts/g.sh <tag> [--gpus N] <remote command>   -> gpurun from the repo root, log under gpurun_out/call_<tag>.log
cd /root/repo || exit 1
tag=$1; shift
extra=""
if [ "$1" = "--gpus" ]; then extra="--gpus $2"; shift 2; fi
mkdir -p gpurun_out
/usr/local/graft/bin/gpurun $extra --timeout ${GTIMEOUT:-1500} -- "$*" > gpurun_out/call_$tag.log 2>&1
tail -2 gpurun_out/call_$tag.log
