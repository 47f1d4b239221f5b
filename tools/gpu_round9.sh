#!/bin/bash
# 2-GPU pass: the driver's torchrun launch of bench.py (IEKF replicas + loop-ICP pairs sharded) and the standalone loop workload.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv,noheader > gpurun_out/gpu_n2.txt 2>&1
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N exit $?"; cut -c1-300 gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --workload loop --steps 3 --warmup 1 > gpurun_out/bench_loop_n$N.json 2> gpurun_out/bench_loop_n$N.err; echo "loop N=$N exit $?"; cut -c1-600 gpurun_out/bench_loop_n$N.json; tail -3 gpurun_out/bench_loop_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref N=$N exit $?"; cut -c1-200 gpurun_out/bench_ref_n$N.json
