#!/bin/bash
# A/B of the view-parallel reduction variants at N GPUs (default: all visible):
#   early  : zero + rank barrier during the forward on a side stream (default)  vs  late (at backward start)
#   bands  : tile-row bands of the blend backward pipelined against the multicast adds
#   same   : every rank renders the same camera (separates camera imbalance from communication cost)
N=${1:-$(nvidia-smi -L | wc -l)}
mkdir -p gpurun_out
run() {  # tag, env..., -- extra bench args
  local tag=$1; shift
  local envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port $((29700 + RANDOM % 200)) bench.py --gpus $N --steps 20 --warmup 5 --no-e2e "$@" 2>&1 | tail -1 > gpurun_out/var_${N}_$tag.json
  python - <<P
import json
try:
    d = json.loads(open("gpurun_out/var_${N}_$tag.json").read())
    print("$tag", d["value"], d["ms_per_step"])
except Exception as e:
    print("$tag", "FAILED", e, open("gpurun_out/var_${N}_$tag.json").read()[-300:])
P
}
#   factored: 64 B / Gaussian exchange + local expand (default)  vs  full 236 B / Gaussian multicast reduction
run factored       GSR_REDUCE_FACTORED=1 --
run factored_late  GSR_REDUCE_FACTORED=1 GSR_REDUCE_EARLY=0 --
run factored_same  GSR_REDUCE_FACTORED=1 -- --yaw-step 0
run full_b4        GSR_REDUCE_FACTORED=0 GSR_REDUCE_BANDS=4 --
run full_b1        GSR_REDUCE_FACTORED=0 GSR_REDUCE_BANDS=1 --
run nccl           GSR_REDUCE_FACTORED=0 -- --reduce nccl
