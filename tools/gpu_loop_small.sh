#!/bin/bash
# per-rank share of the 8-GPU loop batch (32 pairs) on one GPU under the ncu time-only pass: which kernels do not shrink
mkdir -p gpurun_out
LOOP_CONTEXTS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_loop32.csv python bench.py --workload loop --pairs 32 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_loop32.log 2>&1; echo "ncu exit $?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_loop32.csv')) if len(r)>5]
hdr=[i for i,r in enumerate(rows) if 'Kernel Name' in r][0]
H=rows[hdr]; ki=H.index('Kernel Name'); vi=H.index('Metric Value')
seq=[(r[ki].split('(')[0].split('::')[-1][:30], float(r[vi].replace(',',''))/1e3) for r in rows[hdr+1:] if r[vi].replace(',','').replace('.','').isdigit()]
# first batch only: up to the first k_icp_finish
out=[]
for n,t in seq:
    out.append((n,t))
    if n.startswith('k_icp_finish'): break
tot=sum(t for _,t in out)
print('kernels', len(out), 'total us', round(tot,1))
for n,t in out: print(f"  {n:32s} {t:9.1f}")
PY
