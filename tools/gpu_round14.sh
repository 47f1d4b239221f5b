#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_loop.py -m gpu -x -q > gpurun_out/pytest_gpu_loop.log 2>&1; echo "pytest loop exit $?"; tail -3 gpurun_out/pytest_gpu_loop.log
for o in morton hilbert; do
LOOP_TREE_ORDER=$o timeout 900 python bench.py --workload loop --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_loop_$o.json 2> gpurun_out/bench_loop_$o.err; echo "loop $o exit $?"; tail -3 gpurun_out/bench_loop_$o.err
python - <<PY
import json
l=json.loads(open('gpurun_out/bench_loop_$o.json').read().strip().splitlines()[-1])
print('$o', l['value'], l['ms_per_batch'], l['rank0']['kernel_ms'])
PY
done
