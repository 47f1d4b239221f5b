#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/map_diff.py > gpurun_out/map_diff.log 2>&1; echo "map_diff exit $?"; tail -40 gpurun_out/map_diff.log
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
