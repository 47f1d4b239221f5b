#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_loop.py tests/test_imu.py -m gpu -x -q > gpurun_out/pytest_gpu_loop.log 2>&1; echo "pytest loop exit $?"; tail -3 gpurun_out/pytest_gpu_loop.log
timeout 900 python bench.py --workload loop --steps 3 --warmup 1 > gpurun_out/bench_loop_n1.json 2> gpurun_out/bench_loop_n1.err; echo "loop exit $?"; tail -3 gpurun_out/bench_loop_n1.err
python - <<PY
import json
l=json.loads(open('gpurun_out/bench_loop_n1.json').read().strip().splitlines()[-1])
print(l['value'], l['ms_per_batch'], l['rank0']['kernel_ms'], l.get('cpu_baseline'))
PY
