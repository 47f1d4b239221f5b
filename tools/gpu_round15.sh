#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_lio.py -m gpu -x -q > gpurun_out/pytest_gpu_lio.log 2>&1; echo "pytest lio exit $?"; tail -3 gpurun_out/pytest_gpu_lio.log
for o in morton hilbert; do
LIO_TILE_ORDER=$o timeout 900 python bench.py --no-loop --no-cpu-baseline > gpurun_out/bench_tile_$o.json 2> gpurun_out/bench_tile_$o.err; echo "bench $o exit $?"; tail -3 gpurun_out/bench_tile_$o.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_tile_$o.json').read().strip().splitlines()[-1])
print('$o', d['value'], d['e2e']['value'], d['kernel_ms'], d['knn_deferred_fraction'])
PY
done
