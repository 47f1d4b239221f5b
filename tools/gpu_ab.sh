#!/bin/bash
# A/B of the tile-search knobs on the bench workload
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --no-loop --no-configs --no-cpu-baseline --no-parity > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$tag.json").read().strip().splitlines()[-1])
print("$tag", round(d["value"],1), round(d["e2e"]["value"],1), {k: (round(v["mean_ms"],4), v["launches"]) for k,v in d["kernel_ms"].items()}, {k: round(v,4) for k,v in d["knn_deferred_fraction"].items()})
PY
}
for spec in "$@"; do
  tag=$(echo "$spec" | tr ' =' '__')
  run "$tag" $spec
done
