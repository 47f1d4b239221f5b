#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-loop > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print(d['value'], d['e2e']['value'], d['ms_per_step'], d['kernel_ms'], d['cpu_baseline']['value'])"; tail -3 gpurun_out/bench.err
timeout 900 python bench.py --workload map_growth > gpurun_out/bench_map.json 2> gpurun_out/bench_map.err; echo "map_growth exit $?"; cut -c1-1800 gpurun_out/bench_map.json; tail -3 gpurun_out/bench_map.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches exit $?"
