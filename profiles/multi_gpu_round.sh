#!/bin/bash
# Runs on a multi-GPU box (gpurun --gpus N): correctness of the fused exchange against NCCL at every rank count, the
# weak-scaling sweep of bench.py at C3 (with its reduce_check / exchange stage / phase gaps), and C4 (6M Gaussians,
# 3840x2160, one camera per GPU: BASELINE.json configs[3]) at 1 and N ranks.
# Usage: bash profiles/multi_gpu_round.sh <tag>
TAG=${1:-r02}
NG=$(nvidia-smi -L | wc -l)
OUT=gpurun_out/${TAG}_scaling
mkdir -p $OUT
trun() { local n=$1; shift; timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
        --master-port $((29500 + RANDOM % 400)) "$@"; }
for N in 2 4 8; do
  [ $N -gt $NG ] && break
  trun $N tests/multi_gpu_check.py > $OUT/check_$N.txt 2>&1; echo "check N=$N rc=$? $(grep -h 'fused multimem\|max rel' $OUT/check_$N.txt | tail -2 | tr '\n' ' ')"
done
for WL in C3 C4; do
  timeout 420 python bench.py --steps 30 --warmup 8 --no-cpu-baseline --workload $WL 2> $OUT/${WL}_1.err | tail -1 > $OUT/${WL}_1.json
  for N in 2 4 8; do
    [ $N -gt $NG ] && break
    [ $WL = C4 ] && [ $N -ne $NG ] && continue
    trun $N bench.py --gpus $N --steps 30 --warmup 8 --workload $WL 2> $OUT/${WL}_$N.err | tail -1 > $OUT/${WL}_$N.json
  done
done
# the NCCL fallback at the largest rank count, for the comparison in DESIGN.md
trun $NG bench.py --gpus $NG --steps 30 --warmup 8 --reduce nccl 2> $OUT/C3_${NG}_nccl.err | tail -1 > $OUT/C3_${NG}_nccl.json
python - <<PY
import glob, json
for f in sorted(glob.glob("$OUT/C*.json")):
    try:
        d = json.loads(open(f).read())
    except Exception as e:
        print(f, "FAILED", open(f.replace(".json", ".err")).read()[-400:]); continue
    st = {s["stage"]: s["ms"] for s in d.get("stages", [])}
    print(f.split("/")[-1], "n", d["n_gpus"], "value", d["value"], "ms", d["ms_per_step"], "e2e", d.get("e2e", {}).get("value"),
          "check", (d.get("reduce_check") or {}).get("max_rel_err"), "exchange", st.get("exchange"), "pre_bwd", st.get("preprocess_bwd"),
          "trace", d.get("trace_ms"))
PY
