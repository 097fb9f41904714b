#!/bin/bash
# One --set full capture of a radix pass in the sort microbenchmark (scripts/bench_sort.py), summarised on the box.
# usage: bash scripts/gpu_prof_sort.sh <tag> <kernel regex> [env assignments...]
TAG=$1; RX=$2; shift 2
mkdir -p gpurun_out
env "$@" timeout 120 python scripts/bench_sort.py 2>&1 | tail -1
env "$@" timeout 200 ncu --set full --clock-control none --import-source on -k regex:$RX -s 4 -c 1 -f -o gpurun_out/${TAG} python scripts/bench_sort.py > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
rep=gpurun_out/${TAG}.ncu-rep
if [ -f $rep ]; then
  ncu -i $rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
  python scripts/ncu_show.py < gpurun_out/${TAG}_raw.csv > gpurun_out/${TAG}_summary.txt 2>&1
  python scripts/ncu_lines.py $rep ${NCU_TOP:-60} > gpurun_out/${TAG}_lines.txt 2>&1
  cat gpurun_out/${TAG}_summary.txt
  head -50 gpurun_out/${TAG}_lines.txt
  rm -f $rep
else
  tail -20 gpurun_out/${TAG}_ncu.log
fi
