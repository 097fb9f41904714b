#!/bin/bash
# first GPU pass: box facts, parity tests, small bench with --check, then the full configs[1] bench
mkdir -p gpurun_out
{ nproc; free -g | head -2; df -h /tmp | tail -1; nvidia-smi --query-gpu=name,memory.total --format=csv,noheader; } > gpurun_out/box.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_rc=$?"
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --genomes 100 --pairs 20000 --len-min 100000 --len-max 200000 --steps 2 --warmup 1 --check --cpu-pairs 20000 \
   > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err; echo "small_rc=$?"
tail -5 gpurun_out/bench_small.err; cat gpurun_out/bench_small.json
timeout 1500 python bench.py --steps 3 --warmup 3 --check > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "full_rc=$?"
tail -15 gpurun_out/bench_full.err; cat gpurun_out/bench_full.json
