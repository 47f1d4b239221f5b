#!/bin/bash
# quick 1-GPU check of the IEKF path: lio GPU tests, tile timeline, bench line without the loop section
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lio.py tests/test_imu.py -m gpu -x -q 2>&1 | tail -15
python tools/tile_prof.py > gpurun_out/tile_prof.txt 2>&1; tail -25 gpurun_out/tile_prof.txt
timeout 300 python bench.py --no-loop --no-configs --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; tail -3 gpurun_out/bench_q.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_q.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["e2e_wall"]["value"], {k: round(v["mean_ms"],4) for k,v in d["kernel_ms"].items()}, d["knn_deferred_fraction"], d["parity_at_config"]["all"])
PY
