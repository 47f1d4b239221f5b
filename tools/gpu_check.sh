#!/bin/bash
# quick 1-GPU check: parity tests + IEKF bench line (+ the 20M-map sweep's exactness sample)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-loop --no-configs --no-cpu-baseline > gpurun_out/bench_check.json 2> gpurun_out/bench_check.err; echo "bench exit $?"; tail -3 gpurun_out/bench_check.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_check.json').read().strip().splitlines()[-1])
print('value', d['value'], 'e2e', d['e2e']['value'], d['kernel_ms'], d['knn_deferred_fraction'])
PY
timeout 900 python bench.py --workload knn_sweep > gpurun_out/bench_sweep.json 2> gpurun_out/bench_sweep.err; echo "sweep exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_sweep.json').read().strip().splitlines()[-1])
print('sweep', d['value'], [(s['filter_size_map'], round(s['knn_ms'],4), s['deferred_fraction'], s['exact_on_sample']) for s in d['sweep']])
PY
