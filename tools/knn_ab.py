"""A/B of the two kNN kernels on the bench workload (GPU box): tile (default) vs per-query groups (LIO_KNN_MODE=group).
Checks that both return identical neighbour ids / counts and prints the per-launch device times."""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("fast-lio-sam_gps_b200")
synth = importlib.import_module("fast-lio-sam_gps_b200.synth")
n_map = int(os.environ.get("N_MAP", 2_000_000))
w = bench.make_iekf_workload(synth, 131072, n_map, 2)
out = {}
ref = None
for mode in ("group", "tile"):
    os.environ["LIO_KNN_MODE"] = mode
    g = pkg.LioContext(0.5, max_points=n_map + (1 << 20), max_scan=131072)
    g.map_build(w["map"])
    res = []
    for j in range(2):
        g.scan_upload(w["scans"][j])
        g.h_share(w["starts"][j], True)  # warm-up (and the sort, for the tile path)
        g.timing(True)
        g.timing_reset()
        for _ in range(10):
            H, z, n_eff, _ = g.h_share(w["starts"][j], True)
        ms, n = g.timing_get("knn")
        g.timing(False)
        o = g.scan_get()
        res.append((o["nn_ids"].copy(), o["nn_cnt"].copy(), H.copy(), n_eff))
        out[f"{mode}_scan{j}"] = {"knn_ms": ms, "launches": n, "n_eff": int(n_eff)}
    out[f"{mode}_fallback_queries"] = g.knn_fallback_queries()
    if ref is None:
        ref = res
    else:
        for a, b in zip(ref, res):
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), "tile and group kNN disagree"
            assert np.array_equal(a[2], b[2]) and a[3] == b[3]
        out["identical"] = True
    g.close()
print(json.dumps(out, indent=1))
