#!/bin/bash
# 1-GPU pass: parity tests, bench line, launch list of the loop-closure batch (which kernels own the pre-ICP time).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_full.json; tail -3 gpurun_out/bench_full.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_full.json').read().strip().splitlines()[-1])
print('kernel_ms', d['kernel_ms']); print('e2e', d['e2e']['value'], 'defer', d['knn_deferred_fraction'])
l=d['loop_icp']; print('loop', l['value'], l['ms_per_batch'], l['rank0']['kernel_ms'], l.get('cpu_baseline'))
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_loop.csv python bench.py --workload loop --pairs 256 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_loop_launch.log 2>&1; echo "ncu loop launches exit $?"
