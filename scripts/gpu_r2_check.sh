#!/bin/bash
# usage: bash scripts/gpu_r2_check.sh <tag> [bench args]: GPU parity tests, then bench (check against ref_harness on)
TAG=${1:-r2}; shift
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest_rc=$?"
tail -5 gpurun_out/${TAG}_pytest.log
timeout 1500 python bench.py --steps 5 --warmup 3 "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench_rc=$?"
tail -4 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json"))
print("value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"], "check", d.get("check"))
print("cpu_baseline", d.get("cpu_baseline"))
for k,v in d["kernels"].items():
    print(f"  {k:28s} {v['ms']/v['launches']:9.3f} ms x{v['launches']}  {v['gbs']} GB/s")
print(d["phases_ms"], d["counts"])
PY
