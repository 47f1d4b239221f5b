"""The step in front of the hot path: IMU forward propagation (esekfom::predict, host FP64) and
UndistortPcl's per-point back-compensation (CUDA).  Checker: oracle/orc_imu.cpp.  Tolerances: predict 1e-12 (same FP64
formulas, different summation structure); undistorted points 1 ulp of f32 at scan range (libm sin/cos may differ in the
last bit between glibc and CUDA before the f32 rounding)."""
import numpy as np
import pytest


def _setup(synth):
    R = synth.rot_from_rpy(0.02, -0.01, 0.4)
    s0 = synth.make_state(R, np.array([3.0, -2.0, 1.8]))
    s0[24:27] = [4.0, 0.5, -0.1]  # vel
    s0[27:30] = [0.002, -0.001, 0.003]  # bg
    s0[30:33] = [0.02, 0.01, -0.03]  # ba
    return s0, synth.default_cov(), synth.imu_noise_cov()


def test_predict_matches_oracle(pkg, orc, synth):
    s0, P0, Q = _setup(synth)
    dt, gyro, acc = synth.make_imu_segment(s0)
    gs, gP, os_, oP = s0, P0, s0, P0
    for k in range(len(gyro)):
        gs, gP = pkg.imu_predict(gs, gP, dt, acc[k], gyro[k], Q)
        os_, oP = orc.imu_predict(os_, oP, dt, acc[k], gyro[k], Q)
        assert np.abs(gs - os_).max() < 1e-12 and np.abs(gP - oP).max() < 1e-12
    assert np.allclose(gP, gP.T, atol=1e-12) and np.linalg.eigvalsh(gP).min() > 0


def test_predict_integrates_constant_motion(pkg, synth):
    """Noise-free constant body rates: position/velocity/rotation follow the closed form; biases, extrinsics, gravity stay."""
    s0, P0, Q = _setup(synth)
    s0[27:33] = 0.0
    R0, g = s0[3:12].reshape(3, 3), s0[33:36]
    omega = np.array([0.0, 0.0, 0.3])
    n, dt = 200, 0.0005
    s = s0
    for k in range(n):
        Rk = s[3:12].reshape(3, 3)
        a_w = np.array([0.5, 0.0, 0.0])  # constant world acceleration
        s, _ = pkg.imu_predict(s, P0, dt, Rk.T @ (a_w - g), omega, Q)
    T = n * dt
    assert np.abs(s[3:12].reshape(3, 3) - R0 @ synth.so3_exp(omega * T)).max() < 1e-9
    assert np.abs(s[24:27] - (s0[24:27] + a_w * T)).max() < 1e-9
    assert np.abs(s[0:3] - (s0[0:3] + s0[24:27] * T + 0.5 * a_w * T * T)).max() < 1e-3  # Euler position step
    assert np.array_equal(s[12:24], s0[12:24]) and np.array_equal(s[27:36], s0[27:36])


def test_predict_small_rate_and_zero_dt(pkg, orc, synth):
    s0, P0, Q = _setup(synth)
    for dt, gy in ((0.0, [0.1, 0.2, 0.3]), (0.005, list(s0[27:30]))):  # dt = 0; omega - bg = 0 exactly
        a, b = pkg.imu_predict(s0, P0, dt, [0.1, 0.2, 9.7], gy, Q), orc.imu_predict(s0, P0, dt, [0.1, 0.2, 9.7], gy, Q)
        assert np.abs(a[0] - b[0]).max() < 1e-12 and np.abs(a[1] - b[1]).max() < 1e-12


def test_undistort_oracle_is_identity_without_motion(orc, synth):
    """A platform at rest (zero rates, zero velocity, gravity-cancelling accel): compensation must not move points."""
    s0 = synth.make_state(np.eye(3), np.zeros(3))
    rows = np.array([np.concatenate([[0.01 * k], np.zeros(3), np.zeros(3), np.zeros(3), s0[0:3], s0[3:12]]) for k in range(11)])
    rng = np.random.default_rng(1)
    pts = rng.uniform(-30, 30, (2000, 3)).astype(np.float32)
    t = np.sort(rng.uniform(0, 100, 2000)).astype(np.float32)
    out = orc.undistort(pts, t, rows, s0)
    assert np.abs(out - pts).max() < 1e-5


@pytest.mark.gpu
def test_undistort_parity(pkg, orc, synth):
    s0, P0, Q = _setup(synth)
    dt, gyro, acc = synth.make_imu_segment(s0, n_imu=20)
    rows, s_end, _ = synth.propagate_poses(orc.imu_predict, s0, P0, dt, gyro, acc, Q)
    rng = np.random.default_rng(3)
    n = 50000
    pts = (rng.normal(0, 1, (n, 3)) * [30, 30, 4]).astype(np.float32)
    t = np.sort(rng.uniform(0, 100.0, n)).astype(np.float32)
    t[:3] = 0.0  # points at the scan start are left alone (time not above the first pose)
    t[-2:] = 103.0  # later than the last IMU pose: extrapolated from the last segment
    want = orc.undistort(pts, t, rows, s_end)
    g = pkg.LioContext(0.5, max_points=1 << 12, max_scan=1 << 16)
    g.scan_upload(pts)
    g.scan_undistort(t, rows, s_end)
    got = g.scan_body()
    assert np.array_equal(got[:3], pts[:3])
    ulp = np.spacing(np.abs(want).astype(np.float32)).max()
    assert np.abs(got - want).max() <= ulp
    assert (got != want).mean() < 1e-3  # almost always bit-identical
    # order independence: the kernel does not need the cloud sorted by time
    perm = rng.permutation(n)
    g.scan_upload(pts[perm])
    g.scan_undistort(t[perm], rows, s_end)
    assert np.array_equal(g.scan_body(), got[perm])
    # error paths
    with pytest.raises(pkg.LioError):
        g.scan_undistort(t[perm], rows[:1], s_end)
    g.close()


@pytest.mark.gpu
def test_undistort_removes_motion_distortion(pkg, orc, synth):
    """Static world points observed from a moving sensor at their own timestamps: after compensation with the product's
    own propagated poses every point matches the end-of-scan observation to IMU-integration accuracy."""
    s0, P0, Q = _setup(synth)
    s0[27:33] = 0.0
    dt, gyro, acc = synth.make_imu_segment(s0, n_imu=40)
    gyro[:] = gyro.mean(0)
    acc[:] = acc.mean(0)
    rows, s_end, _ = synth.propagate_poses(pkg.imu_predict, s0, P0, dt, gyro, acc, Q)
    rng = np.random.default_rng(4)
    n = 20000
    world = (rng.normal(0, 1, (n, 3)) * [25, 25, 3] + s0[0:3]).astype(np.float64)
    t = np.sort(rng.uniform(0.01, 99.99, n)).astype(np.float32)
    # sensor pose at each point's time, interpolated the way UndistortPcl models it (pose k + constant rate / accel)
    k = np.minimum((t / 1000.0 / dt).astype(int), len(rows) - 2)
    d = t / 1000.0 - rows[k, 0]
    offR, offT = s_end[12:21].reshape(3, 3), s_end[21:24]
    body = np.zeros((n, 3), np.float32)
    for i in range(n):
        Ri = rows[k[i], 13:22].reshape(3, 3) @ synth.so3_exp(rows[k[i] + 1, 4:7] * d[i])
        pi = rows[k[i], 10:13] + rows[k[i], 7:10] * d[i] + 0.5 * rows[k[i] + 1, 1:4] * d[i] ** 2
        body[i] = offR.T @ (Ri.T @ (world[i] - pi) - offT)
    Re, pe = s_end[3:12].reshape(3, 3), s_end[0:3]
    want = (offR.T @ (Re.T @ (world - pe).T - offT[:, None])).T
    g = pkg.LioContext(0.5, max_points=1 << 12, max_scan=1 << 16)
    g.scan_upload(body)
    g.scan_undistort(t, rows, s_end)
    got = g.scan_body()
    assert np.abs(body - want).max() > 0.05  # the distortion was real
    assert np.abs(got - want).max() < 2e-4
    g.close()
