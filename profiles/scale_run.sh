#!/bin/bash
# Scaling sweep on one multi-GPU box: bench.py at N = 1, 2, 4, 8 (as many as visible), fused and NCCL reduction.
NG=$(nvidia-smi -L | wc -l)
mkdir -p gpurun_out
for N in 1 2 4 8; do
  [ $N -gt $NG ] && break
  if [ $N -eq 1 ]; then
    timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/scale_ours_$N.json
  else
    for R in fused nccl; do
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) \
        bench.py --gpus $N --steps 20 --warmup 5 --reduce $R 2>&1 | tail -1 > gpurun_out/scale_ours_${N}_$R.json
    done
  fi
done
for f in gpurun_out/scale_ours_*.json; do echo "== $f"; python profiles/show_stages.py < $f; done
