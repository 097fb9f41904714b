#!/bin/bash
# Steady state of `metabuli classify`: the configs[1] sample files concatenated K times (K M pairs), one GPU and N GPUs.
# usage: bash scripts/gpu_cli_big.sh <tag> <K> <N>
TAG=$1; K=${2:-8}; N=${3:-1}
mkdir -p gpurun_out
W=/tmp/mb_bench/g10000_l200000-500000_s2
python bench.py --steps 1 --warmup 1 --no-fastq --no-check > gpurun_out/${TAG}_prep.json 2> gpurun_out/${TAG}_prep.err || { tail -5 gpurun_out/${TAG}_prep.err; exit 1; }
for m in 1 2; do : > /tmp/big_$m.fq; for i in $(seq $K); do cat $W/sample1000000_$m.fq >> /tmp/big_$m.fq; done; done
ls -la /tmp/big_1.fq
OUT=/tmp/mb_cli_out; rm -rf $OUT; mkdir -p $OUT
CLI=metabuli_work_b200/metabuli
for run in 1 2; do
  MB_TIMING=1 $CLI classify /tmp/big_1.fq /tmp/big_2.fq $W/db $OUT big > gpurun_out/${TAG}_big.log 2>&1 || tail -3 gpurun_out/${TAG}_big.log
  grep -E "classify\]|Time for processing|GPU time" gpurun_out/${TAG}_big.log
done
for mb in 64 128 256; do
  echo "MB_CHUNK_MB=$mb"; MB_CHUNK_MB=$mb MB_TIMING=1 $CLI classify /tmp/big_1.fq /tmp/big_2.fq $W/db $OUT bigc > gpurun_out/${TAG}_bigc$mb.log 2>&1
  grep -E "classify\]|Time for processing|GPU time" gpurun_out/${TAG}_bigc$mb.log
  cmp $OUT/big_classifications.tsv $OUT/bigc_classifications.tsv && echo same
done
MB_TIMING=1 MB_HOST_PARSE=1 $CLI classify /tmp/big_1.fq /tmp/big_2.fq $W/db $OUT bigh > gpurun_out/${TAG}_bigh.log 2>&1; grep -E "classify\]|Time for processing|GPU time" gpurun_out/${TAG}_bigh.log
cmp $OUT/big_classifications.tsv $OUT/bigh_classifications.tsv && echo "host-parse == device-parse"
if [ "$N" -gt 1 ]; then
  $CLI classify /tmp/big_1.fq /tmp/big_2.fq $W/db $OUT bigN --gpus $N > gpurun_out/${TAG}_bigN.log 2>&1 || tail -3 gpurun_out/${TAG}_bigN.log
  grep -E "Time for processing" gpurun_out/${TAG}_bigN.log
  cmp $OUT/big_classifications.tsv $OUT/bigN_classifications.tsv && echo "--gpus $N == --gpus 1"
fi
