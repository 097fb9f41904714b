#!/bin/bash
# Round-end evidence on one GPU: default bench line (with cpu_baseline and the full check), reference arm, launch list.
mkdir -p gpurun_out
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference.json 2> gpurun_out/r02_bench_reference.err; tail -c 600 gpurun_out/r02_bench_reference.json
python bench.py > gpurun_out/r02_bench_default.json 2> gpurun_out/r02_bench_default.err; tail -c 300 gpurun_out/r02_bench_default.json
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-check"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 600 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/r02_ncu_launches.log 2>&1
echo "launch list rc=$?"
