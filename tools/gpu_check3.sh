#!/bin/bash
# 1-GPU check: parity tests, bench line (with parity_at_config), launch list, ncu --set full of the search kernels
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-loop > gpurun_out/bench_check.json 2> gpurun_out/bench_check.err; echo "bench exit $?"; tail -3 gpurun_out/bench_check.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_check.json').read().strip().splitlines()[-1])
print('value', d['value'], 'e2e', d['e2e']['value'], 'wall', d.get('e2e_wall',{}).get('value'), 'clocks', d.get('clocks'))
print({k: round(v['mean_ms'],4) for k,v in d['kernel_ms'].items()}, d['knn_deferred_fraction'])
print('parity', d.get('parity_at_config')); print('cpu', d.get('cpu_baseline'))
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_dbg.csv python bench.py --no-loop --no-configs --no-parity --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/ncu_dbg.log 2>&1; echo "ncu exit $?"
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open('gpurun_out/launches_dbg.csv')) if len(r)>5]
hdr=[i for i,r in enumerate(rows) if 'Kernel Name' in r][0]
H=rows[hdr]; ki=H.index('Kernel Name'); vi=H.index('Metric Value')
acc=collections.defaultdict(list)
for r in rows[hdr+1:]:
    try: acc[r[ki][:60]].append(float(r[vi].replace(',','')))
    except: pass
for k,v in sorted(acc.items(), key=lambda kv:-sum(kv[1])): print(f"{k:62s} n={len(v):3d} mean_us={sum(v)/len(v)/1e3:8.2f} total_us={sum(v)/1e3:9.1f}")
PY
if [ -n "$NCU_FULL" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$NCU_FULL" -s 4 -c 2 -o gpurun_out/prof_full python tools/tile_prof.py > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"; ls -la gpurun_out/*.ncu-rep
fi
