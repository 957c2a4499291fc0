#!/bin/bash
# Runs on the GPU box (under gpurun): plain bench run, then the ncu launch list and one
# `--set full` capture of the hot kernels with the same command line.
# Usage: bash profiles/capture.sh <tag> [kernel-regex] [skip] [count]
set -u
TAG=${1:-r01}
RX=${2:-"^(render|preprocess|onesweep|radix|scan|emit|tile_ranges)"}
SKIP=${3:-1}
COUNT=${4:-3}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launch_$TAG.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$RX -s $SKIP -c $COUNT \
    -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full capture rc=$?"
tail -3 gpurun_out/ncu_full_$TAG.log
