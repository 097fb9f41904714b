#!/bin/bash
# Wall clock of `metabuli classify` on the configs[1] files (1M pairs as two FASTQ files, 10k-genome DB directory),
# one GPU and N GPUs, outputs compared.  usage: bash scripts/gpu_cli_wall.sh <tag> <N>
TAG=$1; N=${2:-1}
mkdir -p gpurun_out
W=/tmp/mb_bench/g10000_l200000-500000_s2
python bench.py --steps 1 --warmup 1 --no-fastq > gpurun_out/${TAG}_prep.json 2> gpurun_out/${TAG}_prep.err || { tail -5 gpurun_out/${TAG}_prep.err; exit 1; }
ls -la $W | head -20
OUT=/tmp/mb_cli_out; rm -rf $OUT; mkdir -p $OUT
CLI=metabuli_work_b200/metabuli
for run in 1 2; do
  $CLI classify $W/sample1000000_1.fq $W/sample1000000_2.fq $W/db $OUT g1 > gpurun_out/${TAG}_g1.log 2>&1 || tail -3 gpurun_out/${TAG}_g1.log
  tail -1 gpurun_out/${TAG}_g1.log
done
grep -E "Time for processing|GPU time|Batch budget" gpurun_out/${TAG}_g1.log
MB_TIMING=1 $CLI classify $W/sample1000000_1.fq $W/sample1000000_2.fq $W/db $OUT g1b --batch-reads 200000 > gpurun_out/${TAG}_g1b.log 2>&1; grep -E "db open|Time for processing" gpurun_out/${TAG}_g1b.log
cmp $OUT/g1_classifications.tsv $OUT/g1b_classifications.tsv && cmp $OUT/g1_report.tsv $OUT/g1b_report.tsv && echo "--batch-reads 200000 == one batch"
MB_HOST_PARSE=1 $CLI classify $W/sample1000000_1.fq $W/sample1000000_2.fq $W/db $OUT g1h > gpurun_out/${TAG}_g1h.log 2>&1; grep -E "Time for processing" gpurun_out/${TAG}_g1h.log
cmp $OUT/g1_classifications.tsv $OUT/g1h_classifications.tsv && echo "host-parse == device-parse"
if [ "$N" -gt 1 ]; then
  for run in 1 2; do
    $CLI classify $W/sample1000000_1.fq $W/sample1000000_2.fq $W/db $OUT gN --gpus $N > gpurun_out/${TAG}_gN.log 2>&1 || tail -3 gpurun_out/${TAG}_gN.log
    grep -E "Time for processing" gpurun_out/${TAG}_gN.log
  done
  cmp $OUT/g1_classifications.tsv $OUT/gN_classifications.tsv && cmp $OUT/g1_report.tsv $OUT/gN_report.tsv && echo "--gpus $N == --gpus 1 (byte-identical)"
fi
# the reference harness' per-read results for the same files (written by the bench run above) vs the CLI's TSV
python - <<PY
import os
W="$W"
tsv=[f for f in os.listdir(W) if f.startswith("ref_pairs") and f.endswith(".results.tsv")]
if tsv:
    ref=open(os.path.join(W,tsv[0])).read().splitlines()
    cli=[l for l in open("$OUT/g1_classifications.tsv").read().splitlines() if not l.startswith("#")]
    bad=0
    for r,c in zip(ref,cli):
        rc=r.split("\t"); cc=c.split("\t")
        if rc[0]!=cc[0] or rc[3]!=cc[2]: bad+=1
    print("CLI vs ref_harness (is_classified, taxID):", len(cli), "reads,", bad, "mismatches")
PY
