#!/bin/bash
# usage: bash scripts/gpu_sweep.sh <tag> ENVVAR v1 v2 ...   — bench once per value of a tuning environment variable
TAG=$1; VAR=$2; shift 2
mkdir -p gpurun_out
for v in "$@"; do
  env $VAR=$v timeout 900 python bench.py --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/${TAG}_$v.json 2> gpurun_out/${TAG}_$v.err || { echo "$v failed"; tail -3 gpurun_out/${TAG}_$v.err; continue; }
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_$v.json"))
k=d["kernels"]
print("$VAR=$v", "ms/step", d["ms_per_step"], {n: round(x["ms"]/x["launches"],2) for n,x in k.items() if x["ms"]/x["launches"]>0.5})
PY
done
