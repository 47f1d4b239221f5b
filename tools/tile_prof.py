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
    passes, served, pushed, is_pushed = word & 0xFF, (word >> 8) & 0xFF, (word >> 16) & 0xFF, (word >> 24) & 1
    print(f"scan {j}: tasks {len(p)} (pushed {int(is_pushed.sum())})  cycles mean {cyc.mean():.0f} p50 {np.percentile(cyc, 50):.0f} p90 {np.percentile(cyc, 90):.0f} "
          f"p99 {np.percentile(cyc, 99):.0f} max {cyc.max()}")
    for k in range(0, 8):
        m = passes == k
        if m.any():
            print(f"  passes={k}: {m.sum():5d} tasks  cycles mean {cyc[m].mean():8.0f}  staged/pass {staged[m].sum() / max(1, passes[m].sum()):6.1f} "
                  f"cells/pass {cells[m].sum() / max(1, passes[m].sum()):6.1f}  queries served {served[m].sum()}")
    print(f"  pushed tasks: {int(is_pushed.sum())}, queries served by them {int(served[is_pushed == 1].sum())}, mean served {served[is_pushed == 1].mean() if is_pushed.any() else 0:.1f}; "
          f"tiles that pushed {(pushed[is_pushed == 0] > 0).sum()}")
    if p.shape[1] >= 8:  # timeline: when (ns after the first tile started) and where (SM) each tile ran
        t0 = p[:, 4].astype(np.int64)
        t1 = p[:, 5].astype(np.int64)
        base = t0.min()
        s_ns, e_ns, smid, stc = t0 - base, t1 - base, p[:, 6], p[:, 7].astype(np.int64)
        print(f"  timeline: last start {s_ns.max()} ns, last end {e_ns.max()} ns; start p50 {np.percentile(s_ns, 50):.0f} p90 {np.percentile(s_ns, 90):.0f} "
              f"p99 {np.percentile(s_ns, 99):.0f}; tile duration ns mean {(e_ns - s_ns).mean():.0f} p99 {np.percentile(e_ns - s_ns, 99):.0f} max {(e_ns - s_ns).max()}")
        print(f"  stage share of tile cycles {stc.sum() / cyc.sum():.3f}; SMs used {len(np.unique(smid))}; tiles per SM min {np.bincount(smid).min()} max {np.bincount(smid).max()}")
        for name, mk in (("tiles", is_pushed == 0), ("pushed", is_pushed == 1)):
            if mk.any():
                print(f"  {name}: {int(mk.sum())} tasks, share of all task cycles {cyc[mk].sum() / cyc.sum():.3f}, cycles mean {cyc[mk].mean():.0f}, stage share {stc[mk].sum() / cyc[mk].sum():.3f}, "
                      f"passes/task {passes[mk].mean():.2f}, cells/pass {cells[mk].sum() / max(1, passes[mk].sum()):.1f}, staged/pass {staged[mk].sum() / max(1, passes[mk].sum()):.1f}, "
                      f"served/task {served[mk].mean():.1f}")
        end_sm = np.array([e_ns[smid == k].max() for k in np.unique(smid)])
        print(f"  per-SM last end ns: min {end_sm.min()} p50 {np.percentile(end_sm, 50):.0f} p90 {np.percentile(end_sm, 90):.0f} max {end_sm.max()}")
        alive = [(int(t), int(((s_ns <= t) & (e_ns > t)).sum())) for t in range(0, int(e_ns.max()) + 1, 4000)]
        print("  tiles alive at t (ns):", alive)
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
