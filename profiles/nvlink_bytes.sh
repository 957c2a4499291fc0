#!/bin/bash
# NVLink bytes per step of the view-parallel exchange, from the driver's per-link data counters around a bench run
# (ncu replays kernels, so it cannot run under a multi-rank launch with rank barriers: these counters are the
# measured NVLink traffic).   Usage (gpurun --gpus N): bash profiles/nvlink_bytes.sh <tag> [steps]
TAG=${1:-r02}
STEPS=${2:-200}
NG=$(nvidia-smi -L | wc -l)
OUT=gpurun_out
mkdir -p $OUT
snap() { nvidia-smi nvlink -gt d -i 0 > $OUT/${TAG}_nvlink_$1.txt 2>&1; }
snap before
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29631 \
    bench.py --gpus $NG --steps $STEPS --warmup 8 --no-e2e 2> $OUT/${TAG}_nvlink_bench.err | tail -1 > $OUT/${TAG}_nvlink_bench.json
snap after
python - <<PY
import json, re
def kib(path):
    tx = rx = 0
    for line in open(path):
        m = re.search(r"Data (Tx|Rx): (\d+) KiB", line)
        if m:
            if m.group(1) == "Tx": tx += int(m.group(2))
            else: rx += int(m.group(2))
    return tx, rx
b, a = kib("$OUT/${TAG}_nvlink_before.txt"), kib("$OUT/${TAG}_nvlink_after.txt")
d = json.loads(open("$OUT/${TAG}_nvlink_bench.json").read())
# steps that crossed NVLink between the snapshots: work measurement 0, warm-up 8, timed STEPS, forward-only none,
# reduce_check 1 fused, traced 10
steps = 8 + $STEPS + 1 + 10
out = {"ranks": $NG, "steps_counted": steps, "gpu0_tx_MB_per_step": round((a[0] - b[0]) * 1024 / 1e6 / steps, 2),
       "gpu0_rx_MB_per_step": round((a[1] - b[1]) * 1024 / 1e6 / steps, 2), "P": d["work"]["P"], "V": d["work"]["V"],
       "model_tx_MB_per_step": round(64 * d["work"]["V"] / 1e6, 2), "value": d["value"], "ms_per_step": d["ms_per_step"],
       "what": "sum over GPU 0's links of the driver's NVLink data counters (nvidia-smi nvlink -gt d) around one bench run; "
               "model: 64 B per contributing Gaussian published once through the multicast address (plus barrier words)"}
print(json.dumps(out))
open("$OUT/${TAG}_nvlink.json", "w").write(json.dumps(out))
PY
head -6 $OUT/${TAG}_nvlink_after.txt
