"""Seeded synthetic worlds, maps and LiDAR scans (SURVEY.md §8(d) "scene generator").

Analytic scene: ground plane + axis-aligned boxes on a street grid (planar, so the plane test mostly
passes) + a few vertical cylinders (curved clutter, so it sometimes fails).  Everything is plain numpy,
deterministic for a given seed, and used by tests/, bench.py and __graft_entry__.smoke() to feed the
GPU path and the CPU oracle with identical inputs.
"""
from __future__ import annotations

import numpy as np

SEED = 20261104


def make_scene(half_extent=60.0, pitch=20.0, box_w=9.0, seed=SEED, cyl_per_block=0.5):
    rng = np.random.default_rng(seed)
    cs = np.arange(-half_extent + pitch / 2, half_extent, pitch)
    boxes = []
    cyls = []
    for cx in cs:
        for cy in cs:
            if abs(cx) < pitch / 2 and abs(cy) < pitch / 2:
                continue  # keep the origin block free for the sensor
            w = box_w * rng.uniform(0.7, 1.0)
            d = box_w * rng.uniform(0.7, 1.0)
            h = rng.uniform(6.0, 22.0)
            ox, oy = rng.uniform(-1.0, 1.0, 2)
            boxes.append([cx + ox - w / 2, cy + oy - d / 2, 0.0, cx + ox + w / 2, cy + oy + d / 2, h])
            if rng.uniform() < cyl_per_block:
                cyls.append([cx + pitch / 2 + rng.uniform(-1, 1), cy + rng.uniform(-3, 3), rng.uniform(0.3, 0.8), rng.uniform(3, 8)])
    return dict(half_extent=float(half_extent), boxes=np.asarray(boxes, np.float64).reshape(-1, 6),
                cyls=np.asarray(cyls, np.float64).reshape(-1, 4))


def _lattice(u0, u1, v0, v1, spacing, rng, jitter):
    us = np.arange(u0 + spacing / 2, u1, spacing)
    vs = np.arange(v0 + spacing / 2, v1, spacing)
    U, V = np.meshgrid(us, vs, indexing="ij")
    U = U.ravel() + rng.uniform(-jitter, jitter, U.size)
    V = V.ravel() + rng.uniform(-jitter, jitter, V.size)
    return U, V


def sample_map(scene, spacing=0.5, noise=0.02, seed=SEED + 1, jitter=0.2, max_points=None):
    """Points on every surface of the scene, about one per `spacing` cell, with normal noise."""
    rng = np.random.default_rng(seed)
    L = scene["half_extent"]
    parts = []
    gx, gy = _lattice(-L, L, -L, L, spacing, rng, jitter)
    keep = np.ones(gx.size, bool)
    for b in scene["boxes"]:
        keep &= ~((gx > b[0]) & (gx < b[3]) & (gy > b[1]) & (gy < b[4]))
    gx, gy = gx[keep], gy[keep]
    parts.append(np.stack([gx, gy, rng.normal(0, noise, gx.size)], 1))
    for b in scene["boxes"]:
        x0, y0, z0, x1, y1, z1 = b
        u, v = _lattice(x0, x1, y0, y1, spacing, rng, jitter)  # roof
        parts.append(np.stack([u, v, z1 + rng.normal(0, noise, u.size)], 1))
        for (a0, a1, fixed, axis) in ((x0, x1, y0, 1), (x0, x1, y1, 1), (y0, y1, x0, 0), (y0, y1, x1, 0)):
            u, v = _lattice(a0, a1, z0, z1, spacing, rng, jitter)
            f = fixed + rng.normal(0, noise, u.size)
            parts.append(np.stack([u, f, v], 1) if axis == 1 else np.stack([f, u, v], 1))
    for c in scene["cyls"]:
        cx, cy, r, h = c
        n_ang = max(6, int(2 * np.pi * r / spacing))
        ang, zz = _lattice(0, 2 * np.pi, 0, h, 1.0, rng, 0.0)
        ang = np.repeat(np.linspace(0, 2 * np.pi, n_ang, endpoint=False), max(1, int(h / spacing)))
        zz = np.tile(np.arange(spacing / 2, h, spacing)[: max(1, int(h / spacing))], n_ang)
        rr = r + rng.normal(0, noise, ang.size)
        parts.append(np.stack([cx + rr * np.cos(ang), cy + rr * np.sin(ang), zz], 1))
    pts = np.concatenate(parts, 0).astype(np.float32)
    rng.shuffle(pts, axis=0)
    if max_points is not None and len(pts) > max_points:
        pts = pts[:max_points]
    return np.ascontiguousarray(pts)


def lidar_dirs(kind="vlp16", phase=0.0, n_az=None):
    """Unit ray directions in the sensor frame, ring-major order."""
    if kind == "vlp16":
        el = np.deg2rad(np.linspace(-15, 15, 16))
        az = np.linspace(0, 2 * np.pi, n_az or 1800, endpoint=False)
    elif kind == "os1_128":
        el = np.deg2rad(np.linspace(-22.5, 22.5, 128))
        az = np.linspace(0, 2 * np.pi, n_az or 1024, endpoint=False)
    elif kind == "avia":
        # non-repetitive rosette inside a 70.4 x 77.2 degree field of view
        n = n_az or 24000
        t = np.arange(n) / n
        a = 2 * np.pi * (t * 61.0 + phase)
        b = 2 * np.pi * (t * 37.0 + 0.618 * phase)
        r = np.cos(b * 0.5 + 0.3) * 0.98
        yaw = np.deg2rad(35.2) * r * np.cos(a)
        pitch = np.deg2rad(38.6) * r * np.sin(a)
        d = np.stack([np.cos(pitch) * np.cos(yaw), np.cos(pitch) * np.sin(yaw), np.sin(pitch)], 1)
        return d
    else:
        raise ValueError(kind)
    E, A = np.meshgrid(el, az + phase, indexing="ij")
    return np.stack([np.cos(E) * np.cos(A), np.cos(E) * np.sin(A), np.sin(E)], -1).reshape(-1, 3)


def raycast(scene, origin, dirs, max_range=120.0, min_range=0.5):
    """Distance along each ray to the nearest surface (inf when nothing is hit within max_range)."""
    o = np.asarray(origin, np.float64)
    d = np.asarray(dirs, np.float64)
    n = len(d)
    t = np.full(n, np.inf)
    # ground z = 0
    with np.errstate(divide="ignore", invalid="ignore"):
        tg = -o[2] / d[:, 2]
    L = scene["half_extent"]
    hit = (tg > min_range) & (np.abs(o[0] + tg * d[:, 0]) < L) & (np.abs(o[1] + tg * d[:, 1]) < L)
    t = np.where(hit, np.minimum(t, tg), t)
    boxes = scene["boxes"]
    if len(boxes):
        c = 0.5 * (boxes[:, :3] + boxes[:, 3:])
        rad = 0.5 * np.linalg.norm(boxes[:, 3:] - boxes[:, :3], axis=1)
        near = boxes[np.linalg.norm(c - o, axis=1) - rad < max_range]
        with np.errstate(divide="ignore", invalid="ignore"):
            inv = 1.0 / d
        for b in near:
            t0 = (b[:3] - o) * inv
            t1 = (b[3:] - o) * inv
            tn = np.nanmax(np.minimum(t0, t1), axis=1)
            tf = np.nanmin(np.maximum(t0, t1), axis=1)
            ok = (tf >= tn) & (tn > min_range)
            t = np.where(ok & (tn < t), tn, t)
    cyls = scene["cyls"]
    if len(cyls):  # a cylinder whose axis is farther than max_range + r cannot be hit within max_range (hits beyond are dropped below)
        cyls = cyls[np.hypot(cyls[:, 0] - o[0], cyls[:, 1] - o[1]) - cyls[:, 2] < max_range]
    for cy in cyls:
        cx, cyy, r, h = cy
        ox, oy = o[0] - cx, o[1] - cyy
        a = d[:, 0] ** 2 + d[:, 1] ** 2
        bq = 2 * (ox * d[:, 0] + oy * d[:, 1])
        cq = ox * ox + oy * oy - r * r
        disc = bq * bq - 4 * a * cq
        with np.errstate(divide="ignore", invalid="ignore"):
            tc = (-bq - np.sqrt(np.maximum(disc, 0))) / (2 * a)
        z = o[2] + tc * d[:, 2]
        ok = (disc > 0) & (tc > min_range) & (z > 0) & (z < h)
        t = np.where(ok & (tc < t), tc, t)
    t[t > max_range] = np.inf
    return t


def rot_from_rpy(roll, pitch, yaw):
    cr, sr, cp, sp, cy, sy = np.cos(roll), np.sin(roll), np.cos(pitch), np.sin(pitch), np.cos(yaw), np.sin(yaw)
    return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                     [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                     [-sp, cp * sr, cp * cr]])


def so3_exp(w):
    w = np.asarray(w, float)
    th = np.linalg.norm(w)
    K = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    if th < 1e-12:
        return np.eye(3) + K
    return np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / th ** 2 * (K @ K)


def make_scan(scene, R, t, kind="vlp16", noise=0.02, seed=SEED + 2, phase=0.0, n_points=None, max_range=120.0, n_az=None):
    """Body-frame points (float32, n x 3) of one scan taken at pose (R, t), ring-major order.

    With n_points the azimuth resolution is raised until at least that many rays return, and the
    first n_points returns in ring order... evenly thinned ... are kept, so the scan has exactly n_points.
    """
    rng = np.random.default_rng(seed)
    R = np.asarray(R, float)
    t = np.asarray(t, float)
    mult = 1
    while True:
        dirs_b = lidar_dirs(kind, phase, None if n_az is None and mult == 1 else (n_az or {"vlp16": 1800, "os1_128": 1024, "avia": 24000}[kind]) * mult)
        dist = raycast(scene, t, dirs_b @ R.T, max_range)
        ok = np.isfinite(dist)
        if n_points is None or ok.sum() >= n_points or mult >= 8:
            break
        mult *= 2
    dirs_b = dirs_b[ok]
    dist = dist[ok] + rng.normal(0, noise, ok.sum())
    pts = (dirs_b * dist[:, None]).astype(np.float32)
    if n_points is not None and len(pts) > n_points:
        idx = np.floor(np.arange(n_points) * (len(pts) / n_points)).astype(np.int64)
        pts = pts[idx]
    return np.ascontiguousarray(pts)


def default_grav():
    """Gravity a few mrad off the S2 chart's pole: state_ikfom's S2 basis degenerates EXACTLY at
    (0, 0, -9.809) (MTK S2_Bx takes its fallback branch there), which a real IMU initialisation never hits."""
    g = np.array([0.02, -0.015, -1.0])
    return g / np.linalg.norm(g) * 9.809


def make_state(R, t, grav=None):
    grav = default_grav() if grav is None else grav
    s = np.zeros(36)
    s[0:3] = t
    s[3:12] = np.asarray(R, float).reshape(9)
    s[12:21] = np.eye(3).reshape(9)
    s[33:36] = grav
    return s


def perturb_state(state, dpos=0.05, drot_deg=0.5, seed=SEED + 3):
    rng = np.random.default_rng(seed)
    s = np.array(state, float)
    dp = rng.normal(size=3)
    dp *= dpos / np.linalg.norm(dp)
    dr = rng.normal(size=3)
    dr *= np.deg2rad(drot_deg) / np.linalg.norm(dr)
    s[0:3] += dp
    s[3:12] = (s[3:12].reshape(3, 3) @ so3_exp(dr)).reshape(9)
    return s


def default_cov():
    """P after a short IMU propagation: loose on pose, tight on the rest."""
    P = np.eye(23) * 1e-4
    P[0:3, 0:3] = np.eye(3) * 1e-2
    P[3:6, 3:6] = np.eye(3) * 1e-3
    P[6:12, 6:12] = np.eye(6) * 1e-8
    return P


def street_poses(n, scene, z=1.8, seed=SEED + 4, step=0.5, pitch=20.0):
    """n poses moving along the (free) x-axis street through the origin block."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        x = -0.35 * scene["half_extent"] + step * i
        y = rng.uniform(-0.5, 0.5)
        R = rot_from_rpy(rng.uniform(-0.02, 0.02), rng.uniform(-0.02, 0.02), rng.uniform(-0.1, 0.1) + 0.01 * i)
        out.append((R, np.array([x, y, z])))
    return out


def pose44(R, t):
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = t
    return T


def make_keyframes(scene, n_pass=40, step=1.0, kind="vlp16", n_az=None, max_range=60.0, seed=SEED + 200, drift_t=(0.6, -0.5, 0.1),
                   drift_rpy_deg=(0.5, -0.4, 2.0), walk=0.01, lateral=0.4, z=1.8, start_x=None):
    """Keyframes of an out-and-back drive along the free x-axis street: pass 1 (indices 0..n_pass-1) outward, pass 2
    (n_pass..2*n_pass-1) back over the same ground, so every pass-2 keyframe has loop candidates in pass 1.

    Returns (clouds, poses_est, poses_true, stamps): clouds[i] is an (n, 4) float32 xyzi cloud in the LiDAR frame
    (PosePcd::pcd_, pose_pcd.hpp:39); poses_est are the pose_corrected_eig_ the loop closer sees: exact on pass 1, true pose
    composed with a slowly wandering drift (about drift_t metres / drift_rpy_deg degrees) on pass 2; stamps 1 s apart."""
    rng = np.random.default_rng(seed)
    x0 = -0.3 * scene["half_extent"] if start_x is None else start_x
    clouds, est, true, stamps = [], [], [], []
    D = pose44(rot_from_rpy(*np.deg2rad(drift_rpy_deg)), np.asarray(drift_t, float))
    for i in range(2 * n_pass):
        back = i >= n_pass
        k = (2 * n_pass - 1 - i) if back else i
        x = x0 + step * k
        y = (lateral if back else 0.0) + 0.05 * np.sin(0.3 * k)
        yaw = (np.pi if back else 0.0) + 0.02 * np.sin(0.2 * k)
        R = rot_from_rpy(0.005 * np.sin(0.5 * k), 0.005 * np.cos(0.4 * k), yaw)
        t = np.array([x, y, z])
        pts = make_scan(scene, R, t, kind=kind, n_az=n_az, seed=seed + 1 + i, max_range=max_range)
        inten = rng.uniform(0, 255, (len(pts), 1)).astype(np.float32)
        clouds.append(np.ascontiguousarray(np.concatenate([pts, inten], 1)))
        T = pose44(R, t)
        true.append(T)
        if back:
            D = D @ pose44(so3_exp(rng.normal(0, walk * 0.2, 3)), rng.normal(0, walk, 3))
            est.append(D @ T)
        else:
            est.append(T.copy())
        stamps.append(float(i))
    return clouds, est, true, stamps


# ------------------------------------------------------------------------------------------- IMU
def imu_noise_cov(gyr_cov=0.1, acc_cov=0.1, b_gyr_cov=1e-4, b_acc_cov=1e-4):
    """Process noise Q (12x12) over (gyro, acc, gyro-bias, acc-bias) with the values of kitti.yaml:16-19."""
    return np.diag([gyr_cov] * 3 + [acc_cov] * 3 + [b_gyr_cov] * 3 + [b_acc_cov] * 3).astype(np.float64)


def make_imu_segment(state0, n_imu=20, scan_period=0.1, omega=(0.05, -0.03, 0.4), acc_body=(0.6, -0.2, 0.1), seed=SEED + 300):
    """IMU samples over one scan period for a platform that turns at `omega` (rad/s, body) and accelerates at `acc_body`
    (m/s^2, body, gravity removed): returns (dt, gyro[n,3], acc[n,3]) with acc the specific force the sensor reads at the
    start pose (R^T (a_w - g) is what get_f inverts), plus small white noise."""
    rng = np.random.default_rng(seed)
    R = np.asarray(state0[3:12]).reshape(3, 3)
    g = np.asarray(state0[33:36])
    dt = scan_period / n_imu
    gyro = np.tile(np.asarray(omega, np.float64), (n_imu, 1)) + rng.normal(0, 1e-3, (n_imu, 3))
    acc = np.tile(np.asarray(acc_body, np.float64) - R.T @ g, (n_imu, 1)) + rng.normal(0, 1e-2, (n_imu, 3))
    return dt, gyro, acc


def propagate_poses(predict, state0, P0, dt, gyro, acc, Q):
    """ImuProcess::UndistortPcl's forward pass with `predict` = esekfom::predict (product or oracle): returns the Pose6D
    rows (n+1, 22: offset_time | acc_s | angvel | vel | pos | rot) and the propagated (state, P) at the scan end."""
    s, P = np.array(state0, np.float64), np.array(P0, np.float64)
    rows = [np.concatenate([[0.0], np.zeros(3), np.zeros(3), s[24:27], s[0:3], s[3:12]])]
    for k in range(len(gyro)):
        s, P = predict(s, P, dt, acc[k], gyro[k], Q)
        R = s[3:12].reshape(3, 3)
        angvel = gyro[k] - s[27:30]
        acc_s = R @ (acc[k] - s[30:33]) + s[33:36]
        rows.append(np.concatenate([[dt * (k + 1)], acc_s, angvel, s[24:27], s[0:3], s[3:12]]))
    return np.array(rows), s, P
