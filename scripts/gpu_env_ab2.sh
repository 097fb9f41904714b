#!/bin/bash
# like gpu_env_ab.sh but prints the sort / join kernels. usage: bash scripts/gpu_env_ab2.sh "<bench args>" VAR v1 v2 ...
ARGS=$1; VAR=$2; shift 2
mkdir -p gpurun_out
for v in "$@"; do
  env $VAR=$v timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-fastq --no-check $ARGS > gpurun_out/ab_${VAR}_$v.json 2> gpurun_out/ab_${VAR}_$v.err || { echo "$v failed"; tail -3 gpurun_out/ab_${VAR}_$v.err; }
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/ab_${VAR}_$v.json") if l.startswith("{")][-1])
k=d["kernels"]
print("$VAR=$v", "step", d["ms_per_step"], {n: (round(k[n]["ms"],2), k[n]["launches"]) for n in k if "kmer" in n or "match" in n}, d["counts"]["classified"])
PY
done
