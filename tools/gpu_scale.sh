#!/bin/bash
# loop-ICP batch (256 pairs) sharded over N ranks; one JSON line per run.  usage: gpu_scale.sh N [ENV=VAL ...]
mkdir -p gpurun_out
N=${1:-8}; shift
tag=n$N$(echo "$*" | tr ' =' '__')
env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --workload loop --steps 5 --warmup 1 > gpurun_out/scale_loop_$tag.json 2> gpurun_out/scale_loop_$tag.err; echo "loop N=$N $* exit $?"
python - <<PY
import json
l=json.loads(open("gpurun_out/scale_loop_$tag.json").read().strip().splitlines()[-1])
print("$tag", round(l["value"],1), round(l["ms_per_batch"],2), l.get("per_rank_ms"), l.get("per_rank_mean_iterations"), {k: round(v["total_ms"],2) for k,v in l["rank0"]["kernel_ms"].items()}, "e2e", round(l["e2e"]["value"],1))
PY
