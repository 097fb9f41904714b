#!/bin/bash
# usage: bash scripts/gpu_cfg.sh <tag> <bench args...>: one bench run, summary printed
TAG=$1; shift
mkdir -p gpurun_out
timeout 1700 python bench.py "$@" > gpurun_out/${TAG}.json 2> gpurun_out/${TAG}.err; echo "${TAG} rc=$?"
tail -6 gpurun_out/${TAG}.err
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}.json"))
    print("value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"], "check", d.get("check"))
    print("cpu", (d.get("cpu_baseline") or {}).get("value"), (d.get("cpu_baseline") or {}).get("phases_s"))
    for k,v in d["kernels"].items():
        print(f"  {k:28s} {v['ms']:9.3f} ms x{v['launches']}  {v['gbs']} GB/s")
    print(d["phases_ms"], d["counts"])
    print(d["config"]["workload"])
except Exception as e:
    print("no json", e)
PY
