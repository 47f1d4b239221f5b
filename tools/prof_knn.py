"""Small driver for ncu: a few h_share calls (kNN redone each time) on the bench workload."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("fast-lio-sam_gps_b200")
synth = importlib.import_module("fast-lio-sam_gps_b200.synth")
n_map = int(os.environ.get("N_MAP", 2_000_000))
w = bench.make_iekf_workload(synth, 131072, n_map, 1)
g = pkg.LioContext(0.5, max_points=n_map + (1 << 20), max_scan=131072)
g.map_build(w["map"])
g.scan_upload(w["scans"][0])
for _ in range(int(os.environ.get("CALLS", 4))):
    g.h_share(w["starts"][0], True)
print("fallback", g.knn_fallback_queries())
g.close()
