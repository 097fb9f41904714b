#!/bin/bash
# round 2 first pass: the no-alloc test alone (fresh process), then the whole GPU suite, then the baseline bench
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2a_noalloc_alone.log 2>&1; echo "noalloc_alone_rc=$?"
tail -3 gpurun_out/r2a_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest_rc=$?"
tail -4 gpurun_out/r2a_pytest.log
timeout 1200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench_rc=$?"
tail -3 gpurun_out/r2a_bench.err
cat gpurun_out/r2a_bench.json
