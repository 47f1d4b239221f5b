"""Per-tile cost distribution of the IEKF search on the bench workload (LIO_TILE_PROF=1): python tools/tile_prof.py"""
import importlib
import os
import sys

import numpy as np

os.environ["LIO_TILE_PROF"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("fast-lio-sam_gps_b200")
synth = importlib.import_module("fast-lio-sam_gps_b200.synth")
w = bench.make_iekf_workload(synth, bench.N_SCAN, bench.N_MAP, 2)
g = pkg.LioContext(0.5, max_points=bench.N_MAP + 8 * bench.N_SCAN, max_scan=bench.N_SCAN, max_cells=bench.N_MAP // 2)
g.map_build(w["map"])
for j in range(2):
    g.scan_upload(w["scans"][j])
    for rep in range(3):
        g.h_share(w["starts"][j], True)
    p = g.tile_profile()
    cyc, word, staged, cells = p[:, 0].astype(np.int64), p[:, 1], p[:, 2], p[:, 3]
    passes, deferred, over = word & 0xFF, (word >> 8) & 0xFF, (word >> 16) & 0xFF
    print(f"scan {j}: tiles {len(p)}  cycles mean {cyc.mean():.0f} p50 {np.percentile(cyc, 50):.0f} p90 {np.percentile(cyc, 90):.0f} "
          f"p99 {np.percentile(cyc, 99):.0f} max {cyc.max()}")
    for k in range(0, 8):
        m = passes == k
        if m.any():
            print(f"  passes={k}: {m.sum():5d} tiles  cycles mean {cyc[m].mean():8.0f}  staged/pass {staged[m].sum() / max(1, passes[m].sum()):6.1f} "
                  f"cells/pass {cells[m].sum() / max(1, passes[m].sum()):6.1f}  deferred queries {deferred[m].sum()} (overflow {over[m].sum()})")
    print("  deferred total", deferred.sum(), "tiles with deferrals", (deferred > 0).sum(), "cycle share of tiles with >=3 passes",
          cyc[passes >= 3].sum() / cyc.sum())
    sc = g.scan_get()
    cnt = sc["nn_cnt"]
    print("  nn_cnt histogram", np.bincount(cnt, minlength=6), " sel", int(sc["sel"].sum()))
    np.savez_compressed(os.path.join(ROOT, "gpurun_out", f"tile_prof_{j}.npz"), prof=p)
# map_incremental list sizes
s, P, info = g.iekf_update_map(w["starts"][0], synth.default_cov(), True, trace=True)
print("iters", info["iters"], "n_eff", info["n_eff"], "knn", info["knn"])
g.scan_upload(w["scans"][1])
g.h_share(w["starts"][1], True)
print("map_incremental lists (add, nodown):", g.map_incremental(w["starts"][1], True))
g.close()
