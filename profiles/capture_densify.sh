#!/bin/bash
# Runs on the GPU box (under gpurun): density-control / PLY benchmark, then the ncu launch list and one
# `--set full` capture of its kernels with the same command line.   Usage: bash profiles/capture_densify.sh <tag>
set -u
TAG=${1:-r02}
CMD="python profiles/bench_densify.py"
mkdir -p gpurun_out
$CMD > gpurun_out/densify_$TAG.json 2> gpurun_out/densify_$TAG.log || { echo "bench_densify failed"; tail -5 gpurun_out/densify_$TAG.log; exit 1; }
python profiles/run_c5_densify.py > gpurun_out/c5_densify_$TAG.json 2> gpurun_out/c5_densify_$TAG.log || tail -5 gpurun_out/c5_densify_$TAG.log
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^(densify|ply_rows|reset_opacity)" -c 200 --csv \
    --log-file gpurun_out/launches_densify_$TAG.csv $CMD > /dev/null 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"^(densify_move|densify_classify|densify_dst|ply_rows)" \
    -s 8 -c 6 -f -o gpurun_out/prof_densify_$TAG $CMD > gpurun_out/ncu_full_densify_$TAG.log 2>&1
echo "full capture rc=$?"
