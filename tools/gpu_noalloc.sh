#!/bin/bash
# the allocation-free entry point alone in a fresh process, then after every other GPU test; then the headline bench line
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_zz.log 2>&1; echo "zz alone exit $?"; tail -15 gpurun_out/pytest_zz.log
timeout 400 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -22 gpurun_out/pytest_gpu.log
timeout 300 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench exit $?"; tail -3 gpurun_out/bench_full.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_full.json').read().strip().splitlines()[-1])
print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'wall', round(d['e2e_wall']['value'],1), 'roofline', round(d['roofline']['frac'],4), 'cpu', round(d['cpu_baseline']['value'],2), 'parity', d['parity_at_config']['all'])
print('loop', round(d['loop_icp']['value'],1)); print('configs', json.dumps(d.get('configs'))[-700:])
PY
