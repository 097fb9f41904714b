#!/bin/bash
# usage: bash scripts/gpu_r2_quick.sh <tag> [pytest -k expr]: GPU parity tests (bounded) then a short bench with the reference check
TAG=${1:-q}; KEXPR=${2:-}
mkdir -p gpurun_out
if [ -n "$KEXPR" ]; then
  timeout 300 python -m pytest tests -m gpu -x -q -k "$KEXPR" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest_rc=$?"
else
  timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest_rc=$?"
fi
tail -25 gpurun_out/${TAG}_pytest.log
