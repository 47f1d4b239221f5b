"""Sweep of the tile-kNN knobs on the bench workload (GPU box): device time of the search per call, deferred fractions."""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("fast-lio-sam_gps_b200")
synth = importlib.import_module("fast-lio-sam_gps_b200.synth")
w = bench.make_iekf_workload(synth, 131072, 2_000_000, 2)
for sub in (1, 2, 3):
    for me in (1, 2, 3):
        os.environ["LIO_TILE_MAXSUB"], os.environ["LIO_TILE_MAXE"] = str(sub), str(me)
        g = pkg.LioContext(0.5, max_points=2_000_000 + (1 << 20), max_scan=131072, max_cells=1_000_000)
        g.map_build(w["map"])
        ms_all = []
        for j in range(2):
            g.scan_upload(w["scans"][j])
            g.h_share(w["starts"][j], True)  # warm-up + the once-per-scan sort
            s0 = g.knn_stats_detail()
            g.timing(True)
            g.timing_reset()
            for _ in range(10):
                g.h_share(w["starts"][j], True)
            ms, n = g.timing_get("knn")
            g.timing(False)
            s1 = g.knn_stats_detail()
            ms_all.append(round(ms * 1e3, 1))
        print(json.dumps({"max_sub": sub, "max_e": me, "knn_us": ms_all,
                          "deferred": {k: round((s1[k] - s0[k]) / 10 / 131072, 4) for k in s0}}), flush=True)
        g.close()
