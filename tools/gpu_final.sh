#!/bin/bash
# Full 1-GPU pass: parity tests, every bench workload, reference arm, ncu launch lists and full captures of the top kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/gpu.txt 2>&1; nproc >> gpurun_out/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench exit $?"; tail -3 gpurun_out/bench_full.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_full.json').read().strip().splitlines()[-1])
print('value', d['value'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'], d['roofline']['traffic'], 'cpu', d['cpu_baseline']['value'])
print('kernel_ms', d['kernel_ms'])
l=d['loop_icp']; print('loop', l['value'], l['ms_per_batch'], l['rank0']['kernel_ms'], l['roofline'], l.get('cpu_baseline'))
PY
timeout 600 python bench.py --impl reference --steps 6 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"; cut -c1-160 gpurun_out/bench_ref.json
timeout 900 python bench.py --workload sequence --seq-scans 100 > gpurun_out/bench_seq.json 2> gpurun_out/bench_seq.err; echo "sequence exit $?"; cut -c1-200 gpurun_out/bench_seq.json; tail -2 gpurun_out/bench_seq.err
timeout 900 python bench.py --workload map_growth > gpurun_out/bench_map.json 2> gpurun_out/bench_map.err; echo "map exit $?"; cut -c1-200 gpurun_out/bench_map.json; tail -2 gpurun_out/bench_map.err
timeout 900 python bench.py --workload knn_sweep > gpurun_out/bench_sweep.json 2> gpurun_out/bench_sweep.err; echo "sweep exit $?"; cut -c1-200 gpurun_out/bench_sweep.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_loop.csv python bench.py --workload loop --pairs 256 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_loop_launch.log 2>&1; echo "ncu loop launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_knn_tile|k_knn_left|k_plane_reduce' -s 9 -c 6 -o gpurun_out/top3_full -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_icp_nn' -c 4 -o gpurun_out/icp_nn_full -f python bench.py --workload loop --pairs 64 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_icp.log 2>&1; echo "ncu icp exit $?"
