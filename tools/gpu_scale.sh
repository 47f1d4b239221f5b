#!/bin/bash
# loop-ICP batch (256 pairs) sharded over N ranks; one JSON line per N
mkdir -p gpurun_out
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --workload loop --steps 5 --warmup 1 > gpurun_out/scale_loop_n$N.json 2> gpurun_out/scale_loop_n$N.err; echo "loop N=$N exit $?"; cut -c1-200 gpurun_out/scale_loop_n$N.json; tail -2 gpurun_out/scale_loop_n$N.err
