#!/bin/bash
# usage: bash scripts/gpu_multi.sh <tag> <N> [bench args]   — reference arm + N-GPU arm, as the driver launches them
TAG=$1; N=$2; shift 2
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 "$@" > gpurun_out/${TAG}_ref.json 2> gpurun_out/${TAG}_ref.err; echo "ref_rc=$?"
tail -2 gpurun_out/${TAG}_ref.err; cat gpurun_out/${TAG}_ref.json
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 --check "$@" \
  > gpurun_out/${TAG}_n${N}.json 2> gpurun_out/${TAG}_n${N}.err; echo "multi_rc=$?"
tail -12 gpurun_out/${TAG}_n${N}.err; cat gpurun_out/${TAG}_n${N}.json
