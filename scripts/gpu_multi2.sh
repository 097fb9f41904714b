#!/bin/bash
# usage: bash scripts/gpu_multi2.sh <tag> <N> [exchange modes...]  — N-GPU bench (fused and/or nccl exchange) with the oracle check
TAG=$1; N=$2; shift 2
MODES=${@:-fused}
mkdir -p gpurun_out
for mode in $MODES; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 --check --no-cpu-baseline --exchange $mode \
    > gpurun_out/${TAG}_n${N}_${mode}.json 2> gpurun_out/${TAG}_n${N}_${mode}.err; echo "$mode rc=$?"
  grep -v "^\*\*\*\|OMP_NUM" gpurun_out/${TAG}_n${N}_${mode}.err | tail -4
  python - <<PY
import json
for line in open("gpurun_out/${TAG}_n${N}_${mode}.json"):
    if line.startswith("{"):
        d = json.loads(line)
        print("$mode N=$N value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"], "check", d.get("check"))
        print("  ", d["counts"].get("phases_ms"))
        print("  ", {k: round(v["ms"] / v["launches"], 2) for k, v in d["kernels"].items()})
PY
done
