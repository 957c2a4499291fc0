#!/bin/bash
# Short version of multi_gpu_round.sh: exchange check at the box's rank count and the C3 bench at 2 / 4 / N ranks.
TAG=${1:-r02}
NG=$(nvidia-smi -L | wc -l)
OUT=gpurun_out/${TAG}_scaling
mkdir -p $OUT
trun() { local n=$1; shift; timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
        --master-port $((29500 + RANDOM % 400)) "$@"; }
trun $NG tests/multi_gpu_check.py > $OUT/check_$NG.txt 2>&1; echo "check N=$NG rc=$?"; grep -h "fused multimem" $OUT/check_$NG.txt | tail -1
for N in 2 4 8; do
  [ $N -gt $NG ] && break
  trun $N bench.py --gpus $N --steps 30 --warmup 8 2> $OUT/C3_$N.err | tail -1 > $OUT/C3_$N.json
done
python - <<PY
import glob, json
for f in sorted(glob.glob("$OUT/C*.json")):
    try:
        d = json.loads(open(f).read())
    except Exception as e:
        print(f, "FAILED", open(f.replace(".json", ".err")).read()[-400:]); continue
    st = {s["stage"]: s["ms"] for s in d.get("stages", [])}
    print(f.split("/")[-1], "n", d["n_gpus"], "value", d["value"], "ms", d["ms_per_step"], "e2e", d.get("e2e", {}).get("value"),
          "check", (d.get("reduce_check") or {}).get("max_rel_err"), "exchange", st.get("exchange"), "pre_bwd", st.get("preprocess_bwd"))
PY
