#!/bin/bash
# sort microbenchmark (scripts/bench_sort.py) per library variant gpurun_variants/lib_<name>.so; usage: gpu_sort_variants.sh v1 v2 ...
cp metabuli_work_b200/libmetabuli_b200.so /tmp/lib_keep.so
for v in "$@"; do
  cp gpurun_variants/lib_$v.so metabuli_work_b200/libmetabuli_b200.so
  echo -n "$v: "; MB_RADIX_V2=1 timeout 120 python scripts/bench_sort.py 2>&1 | tail -1
done
cp /tmp/lib_keep.so metabuli_work_b200/libmetabuli_b200.so
echo -n "v1: "; MB_RADIX_V2=0 timeout 120 python scripts/bench_sort.py 2>&1 | tail -1
