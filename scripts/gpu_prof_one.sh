#!/bin/bash
# One --set full capture of one kernel of the bench command, summarised on the box.
# usage: bash scripts/gpu_prof_one.sh <tag> <kernel regex> [launches to skip] [launches to capture] [env assignments...]
TAG=$1; RX=$2; SKIP=${3:-2}; CNT=${4:-1}; shift 4
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-check $BENCH_EXTRA"
env "$@" ncu --set full --clock-control none --import-source on -k regex:$RX -s $SKIP -c $CNT -f -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
rep=gpurun_out/${TAG}.ncu-rep
if [ -f $rep ]; then
  ncu -i $rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
  python scripts/ncu_show.py < gpurun_out/${TAG}_raw.csv > gpurun_out/${TAG}_summary.txt 2>&1
  python scripts/ncu_lines.py $rep ${NCU_TOP:-70} > gpurun_out/${TAG}_lines.txt 2>&1
  cat gpurun_out/${TAG}_summary.txt
  head -45 gpurun_out/${TAG}_lines.txt
  rm -f $rep
else
  tail -20 gpurun_out/${TAG}_ncu.log
fi
