#!/bin/bash
# usage: bash scripts/gpu_variants2.sh "<bench args>" v1 v2 ...: per variant the step time and the scoring kernels
ARGS=$1; shift
mkdir -p gpurun_out
cp metabuli_work_b200/libmetabuli_b200.so /tmp/lib_keep.so
for v in "$@"; do
  cp gpurun_variants/lib_$v.so metabuli_work_b200/libmetabuli_b200.so
  timeout 900 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-check --no-fastq $ARGS > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err || { echo "$v failed"; tail -3 gpurun_out/var_$v.err; }
  python - <<PY
import json
d=json.load(open("gpurun_out/var_$v.json"))
k=d["kernels"]
print("$v", "step", d["ms_per_step"], {n: round(k[n]["ms"],2) for n in k if "score" in n or n in ("k_match","k_sort_segments")}, "classified", d["counts"]["classified"])
PY
done
cp /tmp/lib_keep.so metabuli_work_b200/libmetabuli_b200.so
