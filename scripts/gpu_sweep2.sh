#!/bin/bash
# usage: bash scripts/gpu_sweep2.sh <tag> "ENV1=a ENV2=b" "ENV1=c" ...  — bench once per environment setting
TAG=$1; shift
mkdir -p gpurun_out
i=0
for setting in "$@"; do
  i=$((i+1))
  env $setting timeout 900 python bench.py --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/${TAG}_$i.json 2> gpurun_out/${TAG}_$i.err || { echo "$setting failed"; tail -3 gpurun_out/${TAG}_$i.err; continue; }
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_$i.json"))
k=d["kernels"]
print("$setting", "ms/step", d["ms_per_step"], {n: round(x["ms"]/x["launches"],2) for n,x in k.items() if x["ms"]/x["launches"]>0.5})
PY
done
