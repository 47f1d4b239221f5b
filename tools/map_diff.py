"""Debug: where does the GPU live map differ from the oracle's after one full-size map_incremental?"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from oracle import orc  # noqa: E402

pkg = importlib.import_module("fast-lio-sam_gps_b200")
synth = importlib.import_module("fast-lio-sam_gps_b200.synth")
orc.build()
w = bench.make_iekf_workload(synth, bench.N_SCAN, bench.N_MAP, 1)
g = pkg.LioContext(0.5, max_points=bench.N_MAP + 8 * bench.N_SCAN, max_scan=bench.N_SCAN, max_cells=bench.N_MAP // 2)
o = orc.OracleLio(0.5, nthreads=0)
g.map_build(w["map"])
o.map_build(w["map"])
gx, gid = g.map_export()
ox, oid = o.map_export()
print("after build: sizes", len(gid), len(oid), "equal", np.array_equal(gid, oid) and np.array_equal(gx, ox))
st = w["truths"][0]
g.scan_upload(w["scans"][0])
o.scan_upload(w["scans"][0])
g.h_share(st, True)
o.h_share(st, True)
a = g.map_incremental(st, True)
b = o.map_incremental(st, True)
print("lists gpu", a, "oracle", b)
gx, gid = g.map_export()
ox, oid = o.map_export()
print("sizes", len(gid), len(oid))
gs, os_ = set(gid.tolist()), set(oid.tolist())
only_g, only_o = sorted(gs - os_), sorted(os_ - gs)
print("only gpu", len(only_g), "only oracle", len(only_o))
gmap = dict(zip(gid.tolist(), gx))
omap = dict(zip(oid.tolist(), ox))
base = len(w["map"])
def vox(p):
    return tuple(np.floor(p / np.float32(0.5)).astype(int))
gv, ov = {}, {}
for i, p in zip(gid.tolist(), gx):
    gv.setdefault(vox(p), []).append(i)
for i, p in zip(oid.tolist(), ox):
    ov.setdefault(vox(p), []).append(i)
for i in only_g[:8]:
    p = gmap[i]
    v = vox(p)
    c = (np.array(v) + 0.5) * 0.5
    print("GPU-only id", i, "(batch idx", i - base, ") p", p, "vox", v, "d2c", float(((p - c) ** 2).sum()), "gpu voxel ids", gv.get(v), "oracle voxel ids", ov.get(v),
          [(j, float(((omap[j] - c) ** 2).sum())) for j in ov.get(v, [])])
for i in only_o[:8]:
    p = omap[i]
    v = vox(p)
    c = (np.array(v) + 0.5) * 0.5
    print("ORACLE-only id", i, "(batch idx", i - base, ") p", p, "vox", v, "d2c", float(((p - c) ** 2).sum()), "oracle voxel ids", ov.get(v), "gpu voxel ids", gv.get(v),
          [(j, float(((gmap[j] - c) ** 2).sum())) for j in gv.get(v, [])])
common = sorted(gs & os_)
diffxyz = [i for i in common[:: max(1, len(common) // 200000)] if not np.array_equal(gmap[i], omap[i])]
print("common ids with different xyz (sampled):", len(diffxyz))
