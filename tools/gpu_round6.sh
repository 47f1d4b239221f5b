#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -8 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-loop --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print(d['value'], d['e2e']['value'], d['ms_per_step'], d['kernel_ms'], d.get('gpu_launches'))"; tail -3 gpurun_out/bench.err
LIO_SORT_GRAPH=0 LIO_MAP_GRAPH=0 timeout 600 python bench.py --no-loop --no-cpu-baseline > gpurun_out/bench_nograph.json 2> gpurun_out/bench_nograph.err; echo "bench nograph exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench_nograph.json'));print(d['value'], d['e2e']['value'], d['ms_per_step'], d['kernel_ms'])"
timeout 600 python bench.py --workload map_growth --map-scans 60 > gpurun_out/bench_map.json 2> gpurun_out/bench_map.err; echo "map_growth exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench_map.json'));print(d['value'], d['ms_per_scan'], d['live_set_identical_to_oracle'])"; tail -3 gpurun_out/bench_map.err
