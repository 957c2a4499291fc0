#!/bin/bash
# One 8-GPU box: weak-scaling sweep N = 1, 2, 4, 8 with the default (factored) reduction, then at N = 8 the
# alternatives (same camera on every rank; full 236 B multicast reduction with 4 bands; NCCL all-reduce).
NG=$(nvidia-smi -L | wc -l)
mkdir -p gpurun_out
show() { python - "$1" "$2" <<'P'
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read())
    print(sys.argv[1], d["value"], d["ms_per_step"], d.get("config", {}).get("gradient_reduction"))
except Exception as e:
    print(sys.argv[1], "FAILED", e, open(sys.argv[2]).read()[-400:])
P
}
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e 2>&1 | tail -1 > gpurun_out/g_scale_1.json
show N=1 gpurun_out/g_scale_1.json
run() {  # tag N env... -- args
  local tag=$1 N=$2; shift 2
  local envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port $((29700 + RANDOM % 200)) bench.py --gpus $N --steps 20 --warmup 5 --no-e2e "$@" 2>&1 | tail -1 > gpurun_out/g_$tag.json
  show $tag gpurun_out/g_$tag.json
}
for N in 2 4 8; do
  [ $N -gt $NG ] && break
  run scale_$N $N GSR_REDUCE_FACTORED=1 --
done
[ $NG -ge 8 ] || exit 0
run same_8 8 GSR_REDUCE_FACTORED=1 -- --yaw-step 0
run full_b4_8 8 GSR_REDUCE_FACTORED=0 GSR_REDUCE_BANDS=4 --
run nccl_8 8 GSR_REDUCE_FACTORED=0 -- --reduce nccl
