"""Host helpers around the loop-closure seam: factor ingredients (fast_lio_sam.cpp:699-717), the trajectory writers
(fast_lio_sam.cpp:14-43, 823-866) and the ATE used for north_star's trajectory bar.  CPU only."""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
tj = importlib.import_module("fast-lio-sam_gps_b200.trajectory")


def _pose(rpy, t):
    r, p, y = rpy
    Rx = np.array([[1, 0, 0], [0, np.cos(r), -np.sin(r)], [0, np.sin(r), np.cos(r)]])
    Ry = np.array([[np.cos(p), 0, np.sin(p)], [0, 1, 0], [-np.sin(p), 0, np.cos(p)]])
    Rz = np.array([[np.cos(y), -np.sin(y), 0], [np.sin(y), np.cos(y), 0], [0, 0, 1]])
    T = np.eye(4)
    T[:3, :3] = Rz @ Ry @ Rx
    T[:3, 3] = t
    return T


def test_readable_stamp_format_and_rounding():
    base = 1_700_000_000
    s = tj.readable_stamp(base + 0.1234)
    assert len(s) == 17 and s.endswith("123") and s[:14] == time.strftime("%Y%m%d%H%M%S", time.localtime(base))
    # 0.9996 s rounds to 1000 ms -> carried into the next second with 000 ms
    s = tj.readable_stamp(base + 0.9996)
    assert s.endswith("000") and s[:14] == time.strftime("%Y%m%d%H%M%S", time.localtime(base + 1))
    assert tj.readable_stamp(base + 0.0006).endswith("001") and tj.readable_stamp(base + 0.0004).endswith("000")


def test_loop_factor_matches_the_reference_composition():
    rng = np.random.default_rng(3)
    T_icp = _pose(rng.uniform(-0.05, 0.05, 3), rng.uniform(-1, 1, 3))
    latest, closest = _pose((0.01, -0.02, 1.0), (10, 2, 0.5)), _pose((0.0, 0.01, -2.0), (9, 3, 0.4))
    assert tj.loop_factor({"valid": False, "T": T_icp, "score": 0.3}, latest, closest, 7, 2) is None
    f = tj.loop_factor({"valid": True, "T": T_icp, "score": 0.3}, latest, closest, 7, 2)
    assert (f["key_from"], f["key_to"]) == (7, 2) and np.array_equal(f["variances"], np.full(6, 0.3))
    # pose_from.between(pose_to): pose_from * measured == pose_to, with pose_from = T_icp * latest (order matters)
    assert np.allclose((T_icp @ latest) @ f["measured"], closest, atol=1e-12)
    assert not np.allclose((latest @ T_icp) @ f["measured"], closest, atol=1e-6)


def test_kitti_and_tum_writers(tmp_path):
    poses = [_pose((0.01 * i, -0.02 * i, 0.3 * i), (i, 2 * i, 0.1 * i)) for i in range(5)]
    stamps = [1_700_000_000 + 0.1 * i for i in range(5)]
    kp = tmp_path / "poses_kitti.txt"
    tj.write_kitti(kp, poses)
    back = tj.read_kitti(kp)
    assert back.shape == (5, 4, 4) and np.allclose(back, np.array(poses), rtol=1e-5, atol=1e-5)  # 6 significant digits
    path = tj.write_tum(str(tmp_path), poses, stamps)
    assert os.path.basename(path) == f"slam_poses_{tj.readable_stamp(stamps[0])}_{tj.readable_stamp(stamps[-1])}.txt"
    lines = open(path).read().splitlines()
    assert lines[0] == "#timestamp x y z qx qy qz qw" and len(lines) == 6
    for T, ts, ln in zip(poses, stamps, lines[1:]):
        tok = ln.split()
        assert tok[0] == tj.readable_stamp(ts) and len(tok) == 8
        x, y, z, qx, qy, qz, qw = map(float, tok[1:])
        assert np.allclose([x, y, z], T[:3, 3], rtol=1e-5, atol=1e-5)
        R = np.array([[1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - qz * qw), 2 * (qx * qz + qy * qw)],
                      [2 * (qx * qy + qz * qw), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - qx * qw)],
                      [2 * (qx * qz - qy * qw), 2 * (qy * qz + qx * qw), 1 - 2 * (qx * qx + qy * qy)]])
        assert np.allclose(R, T[:3, :3], atol=2e-5)


def test_ate_is_invariant_to_a_rigid_offset():
    rng = np.random.default_rng(5)
    ref = np.cumsum(rng.normal(0, 1, (200, 3)), 0)
    T = _pose((0.2, -0.1, 1.3), (5, -3, 2))
    est = ref @ T[:3, :3].T + T[:3, 3]
    assert tj.ate(est, ref) < 1e-9 and tj.ate(est, ref, align=False) > 1.0
    noisy = est + rng.normal(0, 0.01, est.shape)
    assert 0.012 < tj.ate(noisy, ref) < 0.022  # sqrt(3) * 0.01
