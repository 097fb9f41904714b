#!/bin/bash
# A/B timing of library variants built into gpurun_variants/lib_<name>.so: bench k_match ms per variant
mkdir -p gpurun_out
cp metabuli_work_b200/libmetabuli_b200.so /tmp/lib_keep.so
for v in "$@"; do
  cp gpurun_variants/lib_$v.so metabuli_work_b200/libmetabuli_b200.so
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-check --no-fastq > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err || { echo "$v failed"; tail -3 gpurun_out/var_$v.err; }
  python - <<PY
import json
d=json.load(open("gpurun_out/var_$v.json"))
k=d["kernels"]
print("$v", "step", d["ms_per_step"], "k_match", round(k["k_match"]["ms"],3), "matches", d["counts"]["matches"], "classified", d["counts"]["classified"])
PY
done
cp /tmp/lib_keep.so metabuli_work_b200/libmetabuli_b200.so
