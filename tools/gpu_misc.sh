#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --workload map_growth --map-scans 2001 --map-check-every 100 > gpurun_out/bench_map_2000.json 2> gpurun_out/bench_map_2000.err; echo "map exit $?"; cut -c1-700 gpurun_out/bench_map_2000.json; tail -2 gpurun_out/bench_map_2000.err
timeout 900 python bench.py --workload sequence --seq-scans 1000 > gpurun_out/bench_seq_1000.json 2> gpurun_out/bench_seq_1000.err; echo "sequence exit $?"; cut -c1-300 gpurun_out/bench_seq_1000.json; tail -2 gpurun_out/bench_seq_1000.err
LIO_KNN_MODE=group N_Q=1000000 timeout 300 python - <<'PY'
import importlib, os, sys, time
sys.path.insert(0, os.getcwd())
import bench_sweep, numpy as np, torch
pkg = importlib.import_module("fast-lio-sam_gps_b200")
m, L = bench_sweep.make_density_map(20_000_000, 0.5)
g = pkg.LioContext(0.5, max_points=21_000_000, max_scan=1_000_000)
g.map_build(m)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for mode in ("spread", "patch"):
    q = bench_sweep.make_queries(m, 1_000_000, mode=mode)
    r = bench_sweep.run_case(g, m, q, 3, flush, torch)
    print("group mode", mode, "ms", round(r["knn_ms"], 3), round(r["knn_ms_l2_flushed"], 3), "Gq/s", round(r["queries_per_s"] / 1e9, 3), r["exact_on_sample"])
PY
