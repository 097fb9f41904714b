#!/bin/bash
# CLI-focused GPU check: the gpu test-suite, then the wall-clock script
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/$1_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/$1_tests.log
bash scripts/gpu_cli_wall.sh $1 ${2:-1}
