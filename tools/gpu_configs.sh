#!/bin/bash
# the full-size standalone workloads of BASELINE.json configs[0] (1000 VLP-16 scans) and configs[2] (2000 Avia scans)
mkdir -p gpurun_out
timeout 1500 python bench.py --workload sequence --seq-scans 1000 > gpurun_out/bench_seq_1000.json 2> gpurun_out/bench_seq_1000.err; echo "sequence exit $?"; cut -c1-600 gpurun_out/bench_seq_1000.json; tail -2 gpurun_out/bench_seq_1000.err
timeout 1500 python bench.py --workload map_growth --map-scans 2001 --map-check-every 100 > gpurun_out/bench_map_2000.json 2> gpurun_out/bench_map_2000.err; echo "map exit $?"; cut -c1-600 gpurun_out/bench_map_2000.json; tail -2 gpurun_out/bench_map_2000.err
