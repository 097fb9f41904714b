#!/bin/bash
# usage: bash scripts/gpu_multi_cfg.sh <tag> <N> [bench args]  — N-GPU arm under torchrun, as the driver launches it
TAG=$1; N=$2; shift 2
mkdir -p gpurun_out
timeout 1700 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 "$@" \
  > gpurun_out/${TAG}_n${N}.json 2> gpurun_out/${TAG}_n${N}.err; echo "multi_rc=$?"
grep -vE "W[0-9]{4}|^\*+$|OMP_NUM" gpurun_out/${TAG}_n${N}.err | tail -12
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_n${N}.json"))
    print("N", d["n_gpus"], "value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"], "check", d.get("check"))
    for k,v in d["kernels"].items():
        print(f"  {k:28s} {v['ms']:9.3f} ms x{v['launches']}  {v['gbs']} GB/s")
    print(d["phases_ms"], d["counts"])
    print(d["config"]["workload"])
except Exception as e:
    print("no json", e)
PY
