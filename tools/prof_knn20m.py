"""Small driver for ncu: 1M queries spread over the 20M-point map of the config-5 sweep (the HBM-bound regime of the search)."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_sweep  # noqa: E402

pkg = importlib.import_module("fast-lio-sam_gps_b200")
n_map, n_q = int(os.environ.get("N_MAP", 20_000_000)), int(os.environ.get("N_Q", 1_000_000))
m, L = bench_sweep.make_density_map(n_map, 0.5)
q = bench_sweep.make_queries(m, n_q, mode="spread" if os.environ.get("SPREAD", "1") == "1" else "patch")
g = pkg.LioContext(0.5, max_points=n_map + (1 << 20), max_scan=n_q)
g.map_build(m)
for _ in range(3):
    ids, d2, _, cnt = g.map_knn(q, 5, 5.0)
print("mean found", cnt.mean(), g.knn_stats_detail())
g.close()
