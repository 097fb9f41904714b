#!/bin/bash
# ncu evidence for one round: launch list of the bench command + one --set full capture per hot kernel.
# The .ncu-rep files are summarised ON THE BOX (raw metrics csv + source-line hot spots) and deleted, except those
# named in KEEP, so that gpurun_out/ stays under the 64 MiB copy-back limit.
# usage: bash scripts/gpu_profile.sh <tag> [keep-regex]     (outputs under gpurun_out/<tag>_*)
TAG=${1:-r01}; KEEP=${2:-k_match$}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-check"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
cat gpurun_out/${TAG}_plain.json
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
for spec in "k_match:^k_match\$:2:1" "k_extract:^k_extract:2:1" "k_radix_pass:^k_radix_pass:18:6" "k_radix_hist:^k_radix_hist:6:2" "k_sort_segments:^k_sort_segments\$:2:1" "k_score_fast:^k_score_fast\$:2:1" "k_score_finish:^k_score_finish\$:2:1" "k_score:^k_score\$:2:1" "k_segment_matches:^k_segment_matches:2:1" "k_match_partition:^k_match_partition:2:1"; do
  IFS=: read name rx skip cnt <<< "$spec"
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -f -o gpurun_out/${TAG}_${name} $CMD > gpurun_out/${TAG}_ncu_${name}.log 2>&1
  echo "$name rc=$?"
  rep=gpurun_out/${TAG}_${name}.ncu-rep
  if [ -f $rep ]; then
    ncu -i $rep --page raw --csv > gpurun_out/${TAG}_${name}_raw.csv 2>/dev/null
    python scripts/ncu_show.py < gpurun_out/${TAG}_${name}_raw.csv > gpurun_out/${TAG}_${name}_summary.txt 2>&1
    python scripts/ncu_lines.py $rep 60 > gpurun_out/${TAG}_${name}_lines.txt 2>&1
    [[ $name =~ $KEEP ]] || rm -f $rep
  fi
done
du -sh gpurun_out/
