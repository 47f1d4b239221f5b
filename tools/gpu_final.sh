#!/bin/bash
# Round-end 1-GPU pass: parity tests, headline bench (with configs), reference arm, sweep, ncu launch lists and full captures.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/gpu.txt 2>&1; nproc >> gpurun_out/gpu.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench exit $?"; tail -3 gpurun_out/bench_full.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_full.json').read().strip().splitlines()[-1])
print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'wall', round(d['e2e_wall']['value'],1), 'roofline', round(d['roofline']['frac'],4), 'cpu', round(d['cpu_baseline']['value'],2), 'parity', d['parity_at_config']['all'])
print('kernel_ms', {k: round(v['mean_ms'],4) for k,v in d['kernel_ms'].items()}, d['knn_deferred_fraction'], d['clocks'])
l=d['loop_icp']; print('loop', round(l['value'],1), round(l['ms_per_batch'],2), {k: round(v['total_ms'],2) for k,v in l['rank0']['kernel_ms'].items()}, l.get('cpu_baseline'))
print('configs', json.dumps(d.get('configs'))[:1500])
PY
timeout 600 python bench.py --impl reference --steps 6 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"; cut -c1-160 gpurun_out/bench_ref.json
timeout 900 python bench.py --workload knn_sweep > gpurun_out/bench_sweep.json 2> gpurun_out/bench_sweep.err; echo "sweep exit $?"; tail -2 gpurun_out/bench_sweep.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_sweep.json').read().strip().splitlines()[-1])
print('probe', d['gather_probe']['GBps_16B_requests'])
for r in d['sweep']+d['sweep_1M']: print(r['filter_size_map'], r['query_set'], r['queries'], 'ms', round(r['knn_ms'],3), round(r['knn_ms_l2_flushed'],3), 'Gq/s', round(r['queries_per_s']/1e9,3), 'frac', round(r['fraction_of_gather_probe_16B'],3), 'sub', round(r['sub_tile_task_fraction'],3), 'pq', round(r['per_query_kernel_fraction'],4), r['exact_on_sample'])
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop --no-configs --no-parity > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_loop.csv python bench.py --workload loop --pairs 256 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_loop_launch.log 2>&1; echo "ncu loop launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_knn_tile|k_knn_left|k_plane_reduce' -s 9 -c 6 -o gpurun_out/top3_full -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop --no-configs --no-parity > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
LOOP_CONTEXTS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_icp_nn' -c 4 -o gpurun_out/icp_nn_full -f python bench.py --workload loop --pairs 64 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_icp.log 2>&1; echo "ncu icp exit $?"
N_MAP=20000000 SPREAD=1 timeout 600 ncu --set full --clock-control none -k regex:'k_knn_tile' -s 1 -c 1 -o gpurun_out/knn_tile_20M -f python tools/prof_knn20m.py > gpurun_out/ncu_20m.log 2>&1; echo "ncu 20M exit $?"
