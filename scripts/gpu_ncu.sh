#!/bin/bash
# usage: bash scripts/gpu_ncu.sh <tag> "<name:regex:skip:count> ..."   — one --set full capture per spec, after a plain run
TAG=$1; SPECS=$2
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
for spec in $SPECS; do
  IFS=: read name rx skip cnt <<< "$spec"
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -f -o gpurun_out/${TAG}_${name} $CMD > gpurun_out/${TAG}_ncu_${name}.log 2>&1
  echo "$name rc=$?"
done
