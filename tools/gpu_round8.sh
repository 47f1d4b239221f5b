#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench exit $?"; cut -c1-400 gpurun_out/bench_full.json; tail -3 gpurun_out/bench_full.err
timeout 600 python bench.py --impl reference --steps 6 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"; cut -c1-200 gpurun_out/bench_ref.json
timeout 900 python bench.py --workload sequence --seq-scans 100 > gpurun_out/bench_seq.json 2> gpurun_out/bench_seq.err; echo "sequence exit $?"; cut -c1-1400 gpurun_out/bench_seq.json; tail -3 gpurun_out/bench_seq.err
timeout 900 python bench.py --workload knn_sweep > gpurun_out/bench_sweep.json 2> gpurun_out/bench_sweep.err; echo "sweep exit $?"; cut -c1-900 gpurun_out/bench_sweep.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_knn_tile|k_knn_left|k_plane_reduce' -s 9 -c 6 -o gpurun_out/top3_full -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
