#!/bin/bash
# Runs on the GPU box (under gpurun): GPU tests, the default bench line, the reference arm, the ncu launch list of a
# short bench run and one `ncu --set full` capture of the path's kernels (each ncu pass only after the same command has
# exited 0 without ncu).  Everything lands in gpurun_out/<tag>_*.
# Usage: bash profiles/gpu_round.sh <tag> [tests|notests] [kernel-regex]
set -u
TAG=${1:-r02}
TESTS=${2:-tests}
RX=${3:-"^(render_|preprocess_|onesweep|radix_|tile_)"}
OUT=gpurun_out
mkdir -p $OUT
if [ "$TESTS" = "tests" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_gputests.log 2>&1
  echo "gpu tests rc=$?"; tail -3 $OUT/${TAG}_gputests.log
fi
timeout 600 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench rc=$?"; cut -c1-600 $OUT/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err
echo "reference rc=$?"; cut -c1-300 $OUT/${TAG}_bench_reference.json
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $SHORT > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv \
    --log-file $OUT/${TAG}_launches.csv $SHORT > $OUT/${TAG}_ncu_launch.log 2>&1
echo "launch list rc=$?"
EXTRA=lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__inst_executed.sum,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active
timeout 300 $SHORT > $OUT/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:"$RX" -s 40 -c 16 \
    -f -o $OUT/${TAG}_prof $SHORT > $OUT/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"; tail -2 $OUT/${TAG}_ncu_full.log
if [ -f $OUT/${TAG}_prof.ncu-rep ]; then
  ncu -i $OUT/${TAG}_prof.ncu-rep --page raw --csv > $OUT/${TAG}_raw.csv 2>/dev/null
  python profiles/ncu_kernels_json.py $OUT/${TAG}_raw.csv > $OUT/${TAG}_kernels.json
  python - <<PY
import json
for k in json.load(open("$OUT/${TAG}_kernels.json")):
    print(k["kernel"][:34].ljust(34), k["us"], "us  dram", k["dram_read_MB"], "+", k["dram_write_MB"], "MB  issue", round(k["issue_active_pct"] or 0, 1), " inst", k["inst"], " regs", k["regs"])
PY
fi
