#!/bin/bash
# loop-ICP: GPU parity tests, then the 256-pair batch with the A/B knobs given as arguments ("A=1" = defaults)
mkdir -p gpurun_out
[ -n "$SKIP_TESTS" ] || timeout 900 python -m pytest tests/test_gpu_loop.py -m gpu -x -q 2>&1 | tail -4
run() { tag=$1; shift; env "$@" timeout 600 python bench.py --workload loop --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/loop_$tag.json 2> gpurun_out/loop_$tag.err; tail -2 gpurun_out/loop_$tag.err; python - <<PY
import json
l=json.loads(open("gpurun_out/loop_$tag.json").read().strip().splitlines()[-1])
print("$tag", round(l["value"],1), round(l["ms_per_batch"],2), {k: round(v["total_ms"],2) for k,v in l["rank0"]["kernel_ms"].items()})
PY
}
for spec in "$@"; do
  tag=$(echo "$spec" | tr ' =' '__')
  run "$tag" $spec
done
