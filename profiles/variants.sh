#!/bin/bash
# A/B of library variants on one GPU: every gaussian_splatting_b200/variants/*.so (built by profiles/build_variants.py
# with -D defines) through the same short bench run; prints ms_per_step and the stage times.
mkdir -p gpurun_out
for so in gaussian_splatting_b200/libgsrast.so gaussian_splatting_b200/variants/*.so; do
  [ -f "$so" ] || continue
  tag=$(basename $so .so)
  GSR_LIB=$PWD/$so timeout 300 python bench.py --steps 30 --warmup 8 --no-cpu-baseline --no-e2e $* > gpurun_out/var_$tag.json 2> gpurun_out/var_$tag.err
  python - <<P
import json
try:
    d = json.loads(open("gpurun_out/var_$tag.json").read())
    print("$tag".ljust(28), d["ms_per_step"], " ".join(f"{s['stage']}={s['ms']}" for s in d["stages"]))
except Exception as e:
    print("$tag", "FAILED", e, open("gpurun_out/var_$tag.err").read()[-300:])
P
done
