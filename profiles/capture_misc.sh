#!/bin/bash
# Runs on the GPU box (under gpurun): `ncu --set full` of the kernels outside the rasterizer's hot path (SSIM, fused
# loss, Adam, the two halves of the exchange on one GPU, density control, PLY rows), each after a plain run of the same
# command; per-launch summaries as JSON.   Usage: bash profiles/capture_misc.sh <tag>
set -u
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
EXTRA=lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__inst_executed.sum,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active
CMD="python profiles/misc_kernels_once.py 2"
timeout 300 $CMD > $OUT/${TAG}_misc_plain.log 2>&1 || { tail -5 $OUT/${TAG}_misc_plain.log; exit 1; }
timeout 900 ncu --set full --metrics $EXTRA --clock-control none --import-source on \
    -k regex:"^(ssim_|loss_|adam_|exchange_|preprocess_bwd)" -s 6 -c 24 -f -o $OUT/${TAG}_misc $CMD > $OUT/${TAG}_misc_ncu.log 2>&1
echo "misc capture rc=$?"
ncu -i $OUT/${TAG}_misc.ncu-rep --page raw --csv > $OUT/${TAG}_misc_raw.csv 2>/dev/null
python profiles/ncu_kernels_json.py $OUT/${TAG}_misc_raw.csv > $OUT/${TAG}_misc_kernels.json
CMD2="python profiles/bench_densify.py"
timeout 300 $CMD2 > $OUT/${TAG}_densify.json 2> $OUT/${TAG}_densify.log &&
timeout 900 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:"^(densify_|ply_rows|reset_opacity)" \
    -s 8 -c 8 -f -o $OUT/${TAG}_densify_prof $CMD2 > $OUT/${TAG}_densify_ncu.log 2>&1
echo "densify capture rc=$?"
ncu -i $OUT/${TAG}_densify_prof.ncu-rep --page raw --csv > $OUT/${TAG}_densify_raw.csv 2>/dev/null
python profiles/ncu_kernels_json.py $OUT/${TAG}_densify_raw.csv > $OUT/${TAG}_densify_kernels.json
python - <<PY
import json
for f in ("$OUT/${TAG}_misc_kernels.json", "$OUT/${TAG}_densify_kernels.json"):
    seen = set()
    for k in json.load(open(f)):
        if k["kernel"] in seen: continue
        seen.add(k["kernel"])
        print(k["kernel"][:44].ljust(44), k["us"], "us dram", k["dram_read_MB"], "+", k["dram_write_MB"], "MB issue", round(k["issue_active_pct"] or 0, 1), "inst", k["inst"], "regs", k["regs"])
PY
