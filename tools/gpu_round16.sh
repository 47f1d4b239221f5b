#!/bin/bash
# N-GPU pass (N = $1): loop-ICP batch sharded over the ranks, then the driver's full bench line.
mkdir -p gpurun_out
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --workload loop --steps 3 --warmup 1 > gpurun_out/bench_loop_n$N.json 2> gpurun_out/bench_loop_n$N.err; echo "loop N=$N exit $?"; cut -c1-400 gpurun_out/bench_loop_n$N.json; tail -2 gpurun_out/bench_loop_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --steps 20 --warmup 4 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N exit $?"; cut -c1-300 gpurun_out/bench_n$N.json; tail -2 gpurun_out/bench_n$N.err
