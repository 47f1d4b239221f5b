#!/bin/bash
# 1-GPU check: parity tests, tile profile, bench line (with parity_at_config)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/tile_prof.py > gpurun_out/tile_prof.log 2>&1; echo "tile_prof exit $?"; cat gpurun_out/tile_prof.log | tail -40
timeout 900 python bench.py --no-loop > gpurun_out/bench_check.json 2> gpurun_out/bench_check.err; echo "bench exit $?"; tail -3 gpurun_out/bench_check.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_check.json').read().strip().splitlines()[-1])
print('value', d['value'], 'e2e', d['e2e']['value'], 'wall', d.get('e2e_wall'), 'clocks', d.get('clocks'))
print({k: round(v['mean_ms'],4) for k,v in d['kernel_ms'].items()}, d['knn_deferred_fraction'])
print('parity', d.get('parity_at_config')); print('cpu', d.get('cpu_baseline'))
PY
