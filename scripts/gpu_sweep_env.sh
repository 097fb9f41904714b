#!/bin/bash
# usage: bash scripts/gpu_sweep_env.sh <tag> "A=1 B=2" "A=2" ...   — bench once per environment assignment set
TAG=$1; shift
mkdir -p gpurun_out
i=0
for e in "$@"; do
  i=$((i+1))
  env $e timeout 900 python bench.py --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/${TAG}_$i.json 2> gpurun_out/${TAG}_$i.err || { echo "$e failed"; tail -3 gpurun_out/${TAG}_$i.err; continue; }
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_$i.json"))
k=d["kernels"]
print("$e", "ms/step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], {n: round(x["ms"]/x["launches"],2) for n,x in k.items() if x["ms"]/x["launches"]>0.5})
PY
done
