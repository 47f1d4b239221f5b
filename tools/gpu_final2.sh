#!/bin/bash
# final pass of the round on the final build: all GPU tests, headline bench, reference arm, loop launch list + ncu capture
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench exit $?"; tail -3 gpurun_out/bench_full.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_full.json').read().strip().splitlines()[-1])
print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'wall', round(d['e2e_wall']['value'],1), 'roofline', round(d['roofline']['frac'],4), 'cpu', round(d['cpu_baseline']['value'],2), 'parity', d['parity_at_config']['all'])
print('kernel_ms', {k: round(v['mean_ms'],4) for k,v in d['kernel_ms'].items()}, d['knn_deferred_fraction'], d['clocks'])
l=d['loop_icp']; print('loop', round(l['value'],1), round(l['ms_per_batch'],2), {k: round(v['total_ms'],2) for k,v in l['rank0']['kernel_ms'].items()}, 'e2e', round(l['e2e']['value'],1), l.get('cpu_baseline'))
print('configs', json.dumps(d.get('configs'))[:1800])
PY
timeout 600 python bench.py --impl reference --steps 6 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"; cut -c1-160 gpurun_out/bench_ref.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_loop.csv python bench.py --workload loop --pairs 256 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_loop_launch.log 2>&1; echo "ncu loop launches exit $?"
LOOP_CONTEXTS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_icp_nn' -c 4 -o gpurun_out/icp_nn_full -f python bench.py --workload loop --pairs 64 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_icp.log 2>&1; echo "ncu icp exit $?"
