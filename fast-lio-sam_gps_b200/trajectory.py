"""Host-side pieces either side of the loop-closure seam (SURVEY.md §8(a-1) last row and §8(f) rows 3-4).  No GPU work
lives here except `PosePcd.from_world_cloud`, which re-expresses the published world-frame cloud in the LiDAR frame with
the same device transform the submap builder uses.

  PosePcd            fast_lio_sam/include/pose_pcd.hpp:7-44      keyframe record; ctor stores the cloud in the LiDAR frame
  loop_factor        fast_lio_sam/src/fast_lio_sam.cpp:699-717   BetweenFactor inputs built from a RegistrationOutput
  readable_stamp     fast_lio_sam/src/fast_lio_sam.cpp:14-43     yyyymmddhhmmssSSS, local time, millisecond rounding
  write_kitti/tum    fast_lio_sam/src/fast_lio_sam.cpp:823-866   poses_kitti.txt (3x4 row-major of pose_corrected) and
                                                                  slam_poses_<start>_<end>.txt (TUM, pose_eig_, RPY->quat)
  ate                north_star's trajectory bar (absolute trajectory error after rigid alignment)
"""
from __future__ import annotations

import os
import time
from dataclasses import dataclass, field

import numpy as np


@dataclass
class PosePcd:
    pcd: np.ndarray  # (n, 4) f32 xyzi in the LiDAR frame
    pose_eig: np.ndarray = field(default_factory=lambda: np.eye(4))
    pose_corrected_eig: np.ndarray = field(default_factory=lambda: np.eye(4))
    timestamp: float = 0.0
    idx: int = 0
    processed: bool = False

    @classmethod
    def from_world_cloud(cls, loop_closure, cloud_world, pose, timestamp, idx):
        """pose_pcd.hpp:38-39: FAST-LIO publishes the registered cloud in the world frame; the keyframe keeps it in the
        LiDAR frame, `transformPcd(tmp, pose.inverse())`.  The transform runs on the device (loop_transform_pcd: f64
        matrix applied to f32 points, like pcl::transformPointCloud with a Matrix4d)."""
        pose = np.asarray(pose, np.float64).reshape(4, 4)
        inv = np.eye(4)
        inv[:3, :3] = pose[:3, :3].T
        inv[:3, 3] = -pose[:3, :3].T @ pose[:3, 3]
        return cls(loop_closure.transform_pcd(cloud_world, inv), pose.copy(), pose.copy(), float(timestamp), int(idx))


def loop_factor(reg_output, latest_pose_corrected, closest_pose_corrected, latest_idx, closest_idx):
    """fast_lio_sam.cpp:699-717.  Returns None when the registration is not valid, else the BetweenFactor's ingredients:
    keys (latest, closest), measured = pose_from.between(pose_to) = pose_from^-1 * pose_to with
    pose_from = T_icp * latest.pose_corrected (order as in the reference), and six equal variances = the fitness score."""
    if not reg_output["valid"]:
        return None
    pose_from = np.asarray(reg_output["T"], np.float64).reshape(4, 4) @ np.asarray(latest_pose_corrected, np.float64).reshape(4, 4)
    pose_to = np.asarray(closest_pose_corrected, np.float64).reshape(4, 4)
    inv = np.eye(4)
    inv[:3, :3] = pose_from[:3, :3].T
    inv[:3, 3] = -pose_from[:3, :3].T @ pose_from[:3, 3]
    return {"key_from": int(latest_idx), "key_to": int(closest_idx), "measured": inv @ pose_to,
            "variances": np.full(6, float(reg_output["score"]))}


def readable_stamp(unix_ts):
    t_sec = int(unix_ts)  # static_cast<time_t>: truncation
    x = (unix_ts - t_sec) * 1000.0
    millis = int(np.copysign(np.floor(abs(x) + 0.5), x))  # std::round: halves away from zero
    if millis == 1000:
        millis, t_sec = 0, t_sec + 1
    return time.strftime("%Y%m%d%H%M%S", time.localtime(t_sec)) + f"{millis:03d}"


def _fmt(v):
    return f"{float(v):.6g}"  # std::ofstream default: 6 significant digits


def _rpy_quat(R):
    """tf::Matrix3x3::getRPY followed by createQuaternionFromRPY (utilities.hpp:60-78) -> (x, y, z, w)."""
    R = np.asarray(R, np.float64)
    if abs(R[2, 0]) >= 1.0:  # gimbal lock branch of getEulerYPR
        yaw = 0.0
        if R[2, 0] < 0:
            pitch, roll = np.pi / 2, np.arctan2(R[0, 1], R[0, 2])
        else:
            pitch, roll = -np.pi / 2, np.arctan2(-R[0, 1], -R[0, 2])
    else:
        pitch = -np.arcsin(R[2, 0])
        cp = np.cos(pitch)
        roll = np.arctan2(R[2, 1] / cp, R[2, 2] / cp)
        yaw = np.arctan2(R[1, 0] / cp, R[0, 0] / cp)
    hy, hp, hr = yaw / 2, pitch / 2, roll / 2
    cy, sy, cp, sp, cr, sr = np.cos(hy), np.sin(hy), np.cos(hp), np.sin(hp), np.cos(hr), np.sin(hr)
    return (sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy)


def write_kitti(path, poses):
    with open(path, "w") as f:
        for T in poses:
            T = np.asarray(T, np.float64).reshape(4, 4)
            f.write(" ".join(_fmt(T[r, c]) for r in range(3) for c in range(4)) + "\n")


def write_tum(directory, poses, stamps):
    """Writes slam_poses_<start>_<end>.txt and returns its path."""
    path = os.path.join(directory, f"slam_poses_{readable_stamp(stamps[0])}_{readable_stamp(stamps[-1])}.txt")
    with open(path, "w") as f:
        f.write("#timestamp x y z qx qy qz qw\n")
        for T, ts in zip(poses, stamps):
            T = np.asarray(T, np.float64).reshape(4, 4)
            q = _rpy_quat(T[:3, :3])
            f.write(" ".join([readable_stamp(ts)] + [_fmt(v) for v in (*T[:3, 3], *q)]) + "\n")
    return path


def read_kitti(path):
    rows = np.loadtxt(path, ndmin=2)
    out = np.tile(np.eye(4), (len(rows), 1, 1))
    out[:, :3, :] = rows.reshape(-1, 3, 4)
    return out


def ate(est_xyz, ref_xyz, align=True):
    """Absolute trajectory error (RMSE of positions) after the least-squares rigid alignment of est onto ref."""
    a, b = np.asarray(est_xyz, np.float64).reshape(-1, 3), np.asarray(ref_xyz, np.float64).reshape(-1, 3)
    if align and len(a) >= 3:
        ca, cb = a.mean(0), b.mean(0)
        U, _, Vt = np.linalg.svd((b - cb).T @ (a - ca))
        D = np.diag([1.0, 1.0, np.sign(np.linalg.det(U @ Vt))])
        R = U @ D @ Vt
        a = (a - ca) @ R.T + cb
    return float(np.sqrt(np.mean(np.sum((a - b) ** 2, axis=1))))
