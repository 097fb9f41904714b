#!/bin/bash
# `metabuli classify` on gzip input: the configs[1] sample files (1M pairs) gzipped; device-parse pipeline (reader threads
# inflate, GPU parses the text) vs the host reader.  usage: bash scripts/gpu_cli_gz.sh <tag>
TAG=$1
mkdir -p gpurun_out
W=/tmp/mb_bench/g10000_l200000-500000_s2
python bench.py --steps 1 --warmup 1 --no-fastq --no-check --no-cpu-baseline > gpurun_out/${TAG}_prep.json 2> gpurun_out/${TAG}_prep.err || { tail -5 gpurun_out/${TAG}_prep.err; exit 1; }
python - <<PY
import sys
sys.path.insert(0, ".")
import bench, numpy as np, torch
sys.argv = ["bench.py"]
a = bench.parse_args()
bases, offsets, lens = bench.prepare_reads(a, torch.device("cuda:0"), 0)
print(bench.write_sample_files(a, bases, offsets, lens, a.pairs))
PY
ls -la $W/*.fq | head
gzip -1 -k -f $W/sample1000000_1.fq $W/sample1000000_2.fq
ls -la $W/*.gz
OUT=/tmp/mb_cli_out; rm -rf $OUT; mkdir -p $OUT
CLI=metabuli_work_b200/metabuli
$CLI classify $W/sample1000000_1.fq $W/sample1000000_2.fq $W/db $OUT plain > gpurun_out/${TAG}_plain.log 2>&1; grep "Time for processing" gpurun_out/${TAG}_plain.log
for run in 1 2; do
  MB_TIMING=1 $CLI classify $W/sample1000000_1.fq.gz $W/sample1000000_2.fq.gz $W/db $OUT gz > gpurun_out/${TAG}_gz.log 2>&1 || tail -3 gpurun_out/${TAG}_gz.log
  grep -E "classify\]|Time for processing" gpurun_out/${TAG}_gz.log
done
MB_TIMING=1 MB_HOST_PARSE=1 $CLI classify $W/sample1000000_1.fq.gz $W/sample1000000_2.fq.gz $W/db $OUT gzh > gpurun_out/${TAG}_gzh.log 2>&1; grep -E "Time for processing" gpurun_out/${TAG}_gzh.log
cmp $OUT/plain_classifications.tsv $OUT/gz_classifications.tsv && cmp $OUT/plain_classifications.tsv $OUT/gzh_classifications.tsv && echo "gz (device parse) == gz (host reader) == plain"
