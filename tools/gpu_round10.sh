#!/bin/bash
# 1-GPU pass: parity tests, bench line, ncu full captures of the search kernels (IEKF kNN + ICP 1-NN).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_full.json; tail -3 gpurun_out/bench_full.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_full.json').read().strip().splitlines()[-1])
print('kernel_ms', d['kernel_ms']); print('e2e', d['e2e']['value'], 'defer', d['knn_deferred_fraction'])
l=d['loop_icp']; print('loop', l['value'], l['ms_per_batch'], l['rank0']['kernel_ms'], l.get('cpu_baseline'))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_knn_tile|k_knn_left' -s 6 -c 2 -o gpurun_out/knn_r01f -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop > gpurun_out/ncu_knn.log 2>&1; echo "ncu knn exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_icp_nn' -s 1 -c 3 -o gpurun_out/icp_nn_r01f -f python bench.py --workload loop --pairs 64 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_icp.log 2>&1; echo "ncu icp exit $?"
