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


if __name__ == "__main__":
    lio_small()
    loop_small()
