#!/bin/bash
# Runs on the GPU box (under gpurun): one `ncu --set full` capture of the kernels matching a regex, after a plain
# run of the same command line has exited 0.   Usage: bash profiles/capture_kernels.sh <tag> <kernel-regex> <skip> <count> [bench args]
set -u
TAG=$1; RX=$2; SKIP=$3; COUNT=$4; shift 4
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e $*"
mkdir -p gpurun_out
timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$RX -s $SKIP -c $COUNT \
    -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full capture rc=$?"
tail -2 gpurun_out/ncu_full_$TAG.log
