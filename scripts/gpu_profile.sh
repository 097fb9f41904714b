#!/bin/bash
# ncu evidence for one round: launch list of the bench command + one --set full capture per hot kernel.
# usage: bash scripts/gpu_profile.sh <tag>      (outputs under gpurun_out/<tag>_*)
TAG=${1:-r01}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
cat gpurun_out/${TAG}_plain.json
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
for spec in "k_match:^k_match\$:2:1" "k_score:^k_score\$:2:1" "k_radix_pass:^k_radix_pass:30:7" "k_extract:^k_extract:4:2"; do
  IFS=: read name rx skip cnt <<< "$spec"
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -f -o gpurun_out/${TAG}_${name} $CMD > gpurun_out/${TAG}_ncu_${name}.log 2>&1
  echo "$name rc=$?"
done
ls -la gpurun_out/
