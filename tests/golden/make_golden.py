"""Regenerates tests/golden/*.npz from the CPU oracle on seeded synthetic inputs.

The reference ships no golden vectors and neither half of its hot path can be built or imported in the
container, so these fixtures pin the ORACLE's behaviour (they catch regressions
and platform drift, e.g. a different libm or compiler contraction setting on the GPU box) — they do not
upgrade parity against the absent reference.   Run:  python tests/golden/make_golden.py
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402

synth = importlib.import_module("fast-lio-sam_gps_b200.synth")
HERE = os.path.dirname(os.path.abspath(__file__))


def lio_small():
    scene = synth.make_scene(half_extent=30.0, seed=synth.SEED + 11)
    m = synth.sample_map(scene, spacing=0.5, seed=synth.SEED + 12, max_points=12000)
    R, t = synth.rot_from_rpy(0.02, -0.01, 0.4), np.array([0.5, -0.3, 1.7])
    scan = synth.make_scan(scene, R, t, kind="vlp16", n_az=120, seed=synth.SEED + 13, max_range=40.0)
    state = synth.perturb_state(synth.make_state(R, t), seed=synth.SEED + 14)
    o = orc.OracleLio(0.5, nthreads=1)
    o.map_build(m)
    o.scan_upload(scan)
    HTH, HTz, n_eff, res = o.h_share(state, True)
    out = o.scan_get()
    P0 = synth.default_cov()
    s, P, it = o.iekf_update(state, P0)
    np.savez_compressed(os.path.join(HERE, "lio_small.npz"), map=m, scan=scan, state=state, nn_ids=out["nn_ids"], sel=out["sel"],
                        plane=out["plane"], n_eff=n_eff, HTH=HTH, HTz=HTz, P0=P0, state_out=s, P_out=P, iters=it)
    print("lio_small: map", m.shape, "scan", scan.shape, "n_eff", n_eff, "iters", it)


def loop_small():
    rng = np.random.default_rng(synth.SEED + 21)
    scene = synth.make_scene(half_extent=30.0, seed=synth.SEED + 11)
    kfs = []
    for i in range(6):
        R, t = synth.rot_from_rpy(0, 0, 0.05 * i), np.array([-2.0 + 0.8 * i, 0.1 * i, 1.7])
        pts = synth.make_scan(scene, R, t, kind="vlp16", n_az=90, seed=synth.SEED + 30 + i, max_range=40.0)
        T = np.eye(4)
        T[:3, :3], T[:3, 3] = R, t
        kfs.append((np.concatenate([pts, rng.uniform(0, 255, (len(pts), 1)).astype(np.float32)], 1), T))
    L = orc.OracleLoop(max_corr_dist=52.5)
    for i, (p, T) in enumerate(kfs):
        Tp = T.copy()
        if i >= 3:  # the "query side" drifted
            Tp[:3, 3] += [0.3, -0.2, 0.05]
        L.add_keyframe(p, Tp, stamp=float(i))
    src, dst = L.set_src_dst(4, 1, 1, 0.3)
    res = L.icp_align(src, dst, want_aligned=True)
    np.savez_compressed(os.path.join(HERE, "loop_small.npz"), **{f"kf{i}": p for i, (p, _) in enumerate(kfs)},
                        poses=np.stack([T for _, T in kfs]), src=src, dst=dst, T=res["T"], score=res["score"],
                        converged=res["converged"], iterations=res["iterations"], state=res["state"], aligned=res["aligned"])
    print("loop_small: src", src.shape, "dst", dst.shape, "iters", res["iterations"], "score", res["score"], "state", res["state"])


def map_small_inputs():
    """Ground patch + three scans: on the patch, in front of it (the far-neighbour branch of map_incremental decides), and
    shifted; one box delete between the second and the third."""
    g = np.arange(-6, 6, 0.5, dtype=np.float32) + 0.13
    gx, gy = np.meshgrid(g, g)
    m = np.stack([gx.ravel(), gy.ravel(), np.full(gx.size, 0.07, np.float32)], 1)
    rng = np.random.default_rng(synth.SEED + 41)
    scans = []
    for k in range(3):
        near = np.stack([rng.uniform(-5, 5, 250), rng.uniform(-5, 5, 250), rng.normal(0.07, 0.02, 250)], 1)
        front = [np.array([8.0 + 2.0 * j + k, 1.1 * j - k, 0.3 * j]) + rng.uniform(0, 0.6, (10, 3)) for j in range(4)]
        far = np.array([[50.0 + k, -44.0, 3.0]])
        scans.append(np.ascontiguousarray(np.concatenate([near] + front + [far]).astype(np.float32)))
    states = np.stack([synth.make_state(synth.rot_from_rpy(0, 0, 0.02 * k), np.array([0.3 * k, 0.1 * k, 0.0])) for k in range(3)])
    boxes = np.array([[-7, -7, -1, -2, 7, 1], [9, -3, -1, 12, 3, 3]], np.float32)  # part of the patch, one front cluster
    return m, scans, states, boxes


def map_small_replay(ctx, m, scans, states, boxes):
    """The same calls on any implementation of the map interface (oracle or the C-ABI view)."""
    ctx.map_build(m)
    counts, deleted = [], -1
    for k, (scan, st) in enumerate(zip(scans, states)):
        ctx.scan_upload(scan)
        ctx.h_share(st, True)
        counts.append(ctx.map_incremental(st, True))
        if k == 1:
            deleted = int(ctx.map_delete_boxes(boxes))
    xyz, ids = ctx.map_export()
    order = np.argsort(ids, kind="stable")
    return np.array(counts, np.int32), deleted, xyz[order], ids[order]


def map_small():
    m, scans, states, boxes = map_small_inputs()
    o = orc.OracleLio(0.5, nthreads=1)
    counts, deleted, xyz, ids = map_small_replay(o, m, scans, states, boxes)
    assert (counts[:, 1] > 0).all() and deleted > 0  # the far-neighbour branch and the delete were exercised
    np.savez_compressed(os.path.join(HERE, "map_small.npz"), map=m, scan0=scans[0], scan1=scans[1], scan2=scans[2], states=states,
                        boxes=boxes, counts=counts, deleted=deleted, xyz=xyz, ids=ids)
    print("map_small: map", m.shape, "counts", counts.tolist(), "deleted", deleted, "final", xyz.shape)


if __name__ == "__main__":
    lio_small()
    loop_small()
    map_small()
