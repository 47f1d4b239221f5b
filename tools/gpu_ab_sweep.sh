#!/bin/bash
mkdir -p gpurun_out
for left in group warp; do for ord in morton hilbert; do
LIO_BATCH_LEFT=$left LIO_TILE_ORDER=$ord timeout 600 python bench.py --workload knn_sweep > gpurun_out/sweep_${left}_${ord}.json 2> gpurun_out/sweep_${left}_${ord}.err; echo "sweep $left $ord exit $?"
python - <<PY
import json
d=json.loads(open('gpurun_out/sweep_${left}_${ord}.json').read().strip().splitlines()[-1])
print('$left $ord', round(d['value']/1e6,1), [(s['filter_size_map'], round(s['knn_ms'],4), round(s['knn_ms_l2_flushed'],4), s['exact_on_sample']) for s in d['sweep']])
PY
done; done
