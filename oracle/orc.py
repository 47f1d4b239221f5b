"""ORACLE — TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (see oracle/README.md).

ctypes view of oracle/liblio_oracle.so, the CPU restatement of the reference's hot path.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module, and only as the checker / the timed CPU baseline — never on the product path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liblio_oracle.so")

_f32p = C.POINTER(C.c_float)
_f64p = C.POINTER(C.c_double)
_i32p = C.POINTER(C.c_int)
_u32p = C.POINTER(C.c_uint32)
_u8p = C.POINTER(C.c_uint8)


def build():
    res = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + res.stdout + res.stderr)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_ctx_create.restype = C.c_void_p
        L.orc_ctx_create.argtypes = [C.c_float, C.c_int]
        L.orc_ctx_destroy.argtypes = [C.c_void_p]
        L.orc_ctx_set.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_double, C.c_float]
        L.orc_max_threads.restype = C.c_int
        L.orc_map_build.argtypes = [C.c_void_p, _f32p, C.c_int]
        L.orc_map_add.restype = C.c_int
        L.orc_map_add.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _i32p]
        L.orc_map_delete_boxes.restype = C.c_int
        L.orc_map_delete_boxes.argtypes = [C.c_void_p, _f32p, C.c_int]
        L.orc_map_size.restype = C.c_int
        L.orc_map_size.argtypes = [C.c_void_p]
        L.orc_map_export.restype = C.c_int
        L.orc_map_export.argtypes = [C.c_void_p, _f32p, _u32p, C.c_int]
        L.orc_map_knn.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, C.c_float, _u32p, _f32p, _f32p, _i32p]
        L.orc_esti_plane.restype = C.c_int
        L.orc_esti_plane.argtypes = [_f32p, C.c_float, _f32p]
        L.orc_scan_set.argtypes = [C.c_void_p, _f32p, C.c_int]
        L.orc_h_share.restype = C.c_int
        L.orc_h_share.argtypes = [C.c_void_p, _f64p, C.c_int, _f64p, _f64p, _f64p, _f64p, _f64p]
        L.orc_scan_get.argtypes = [C.c_void_p, _f32p, _u32p, _f32p, _i32p, _u8p, _f32p]
        L.orc_iekf_update.restype = C.c_int
        L.orc_iekf_update.argtypes = [C.c_void_p, _f64p, _f64p, C.c_double, C.c_int, _f64p, _f64p, _i32p, _i32p]
        L.orc_map_incremental.argtypes = [C.c_void_p, _f64p, C.c_int, _i32p, _i32p]
        L.orc_fov_segment.restype = C.c_int
        L.orc_fov_segment.argtypes = [C.c_void_p, _f64p, _f32p]
        L.orc_boxplus.argtypes = [_f64p, _f64p]
        L.orc_boxminus.argtypes = [_f64p, _f64p, _f64p]
        L.orc_imu_predict.restype = None
        L.orc_imu_predict.argtypes = [_f64p, _f64p, C.c_double, _f64p, _f64p, _f64p]
        L.orc_undistort.restype = None
        L.orc_undistort.argtypes = [_f32p, _f32p, C.c_int, _f64p, C.c_int, _f64p]
        # loop closure
        L.orc_loop_create.restype = C.c_void_p
        L.orc_loop_create.argtypes = [C.c_double, C.c_double, C.c_double, C.c_int, C.c_int]
        L.orc_loop_destroy.argtypes = [C.c_void_p]
        L.orc_loop_add_keyframe.restype = C.c_int
        L.orc_loop_add_keyframe.argtypes = [C.c_void_p, _f32p, C.c_int, _f64p, C.c_double]
        L.orc_loop_set_pose.argtypes = [C.c_void_p, C.c_int, _f64p]
        L.orc_loop_fetch_closest.restype = C.c_int
        L.orc_loop_fetch_closest.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double]
        L.orc_transform_pcd.argtypes = [_f32p, C.c_int, _f64p, _f32p]
        L.orc_pose_inverse.restype = C.c_int
        L.orc_pose_inverse.argtypes = [_f64p, _f64p]
        L.orc_voxelize.restype = C.c_int
        L.orc_voxelize.argtypes = [_f32p, C.c_int, C.c_float, _f32p]
        L.orc_loop_set_src_dst.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, _f32p, _i32p, _f32p, _i32p, C.c_int]
        L.orc_icp_align.argtypes = [C.c_void_p, _f32p, C.c_int, _f32p, C.c_int, _f32p, _f64p, _i32p, _i32p, _i32p, _f32p, _i32p]
        L.orc_loop_verify.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_double, _f32p, _f64p, _i32p, _i32p, _i32p,
                                      _i32p, _i32p]
        L.orc_svd3.argtypes = [_f32p, _f32p, _f32p, _f32p]
        _lib = L
    return _lib


def _p(a, ty):
    return None if a is None else a.ctypes.data_as(ty)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def max_threads():
    """Host threads the CPU legs use: every core this process may run on.  Not omp_get_max_threads(): torchrun exports
    OMP_NUM_THREADS=1 to its workers, which would silently time the OpenMP oracle on one core."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or lib().orc_max_threads())


class OracleLio:
    """CPU restatement of ikd-Tree + h_share_model + esekfom update + map_incremental."""

    def __init__(self, filter_size_map=0.5, nthreads=0, extrinsic_est_en=False, max_sqdist=5.0, cube_side_length=1000.0,
                 det_range=300.0):
        self.L = lib()
        self.h = C.c_void_p(self.L.orc_ctx_create(filter_size_map, nthreads))
        self.L.orc_ctx_set(self.h, int(extrinsic_est_en), max_sqdist, cube_side_length, det_range)
        self.scan_n = 0

    def close(self):
        if self.h:
            self.L.orc_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def map_build(self, xyz):
        xyz = _f32(xyz).reshape(-1, 3)
        self.L.orc_map_build(self.h, _p(xyz, _f32p), len(xyz))

    def map_add(self, xyz, downsample=True):
        xyz = _f32(xyz).reshape(-1, 3)
        net = C.c_int(0)
        calls = self.L.orc_map_add(self.h, _p(xyz, _f32p), len(xyz), int(downsample), C.byref(net))
        return net.value, calls

    def map_delete_boxes(self, boxes):
        boxes = _f32(boxes).reshape(-1, 6)
        return self.L.orc_map_delete_boxes(self.h, _p(boxes, _f32p), len(boxes))

    def map_size(self):
        return self.L.orc_map_size(self.h)

    def map_export(self):
        n = self.map_size()
        xyz = np.zeros((max(n, 1), 3), np.float32)
        ids = np.zeros(max(n, 1), np.uint32)
        tot = self.L.orc_map_export(self.h, _p(xyz, _f32p), _p(ids, _u32p), n)
        return xyz[:tot], ids[:tot]

    def map_knn(self, q, k=5, max_sqdist=5.0):
        q = _f32(q).reshape(-1, 3)
        n = len(q)
        ids = np.zeros((n, k), np.uint32)
        d2 = np.zeros((n, k), np.float32)
        pts = np.zeros((n, k, 3), np.float32)
        cnt = np.zeros(n, np.int32)
        self.L.orc_map_knn(self.h, _p(q, _f32p), n, k, max_sqdist, _p(ids, _u32p), _p(d2, _f32p), _p(pts, _f32p), _p(cnt, _i32p))
        return ids, d2, pts, cnt

    def scan_upload(self, body_xyz):
        body_xyz = _f32(body_xyz).reshape(-1, 3)
        self.scan_n = len(body_xyz)
        self.L.orc_scan_set(self.h, _p(body_xyz, _f32p), len(body_xyz))

    def h_share(self, state, redo_knn=True, want_rows=False):
        state = _f64(state).reshape(36)
        HTH = np.zeros((12, 12))
        HTz = np.zeros(12)
        res = C.c_double(0)
        hx = np.zeros((self.scan_n, 12)) if want_rows else None
        hv = np.zeros(self.scan_n) if want_rows else None
        n_eff = self.L.orc_h_share(self.h, _p(state, _f64p), int(redo_knn), _p(HTH, _f64p), _p(HTz, _f64p), C.byref(res),
                                   _p(hx, _f64p), _p(hv, _f64p))
        if want_rows:
            return HTH, HTz, n_eff, res.value, hx[:n_eff], hv[:n_eff]
        return HTH, HTz, n_eff, res.value

    def scan_get(self):
        n = self.scan_n
        out = dict(world=np.zeros((n, 3), np.float32), nn_ids=np.zeros((n, 5), np.uint32), nn_xyz=np.zeros((n, 5, 3), np.float32),
                   nn_cnt=np.zeros(n, np.int32), sel=np.zeros(n, np.uint8), plane=np.zeros((n, 4), np.float32))
        self.L.orc_scan_get(self.h, _p(out["world"], _f32p), _p(out["nn_ids"], _u32p), _p(out["nn_xyz"], _f32p), _p(out["nn_cnt"], _i32p),
                            _p(out["sel"], _u8p), _p(out["plane"], _f32p))
        return out

    def iekf_update(self, state, P, R=0.001, max_iter=3, limit=None, trace=False):
        state = _f64(state).reshape(36).copy()
        P = _f64(P).reshape(23, 23).copy()
        limit = _f64(np.full(23, 0.001) if limit is None else limit)
        it_states = np.zeros((max_iter + 1, 36))
        n_eff = np.zeros(max_iter + 1, np.int32)
        knn = np.zeros(max_iter + 1, np.int32)
        done = self.L.orc_iekf_update(self.h, _p(state, _f64p), _p(P, _f64p), R, max_iter, _p(limit, _f64p), _p(it_states, _f64p),
                                      _p(n_eff, _i32p), _p(knn, _i32p))
        if trace:
            return state, P, dict(iters=done, states=it_states[:done], n_eff=n_eff[:done], knn=knn[:done])
        return state, P, done

    def map_incremental(self, state, flg_ekf_inited=True):
        state = _f64(state).reshape(36)
        a, b = C.c_int(0), C.c_int(0)
        self.L.orc_map_incremental(self.h, _p(state, _f64p), int(flg_ekf_inited), C.byref(a), C.byref(b))
        return a.value, b.value

    def fov_segment(self, pos_lid):
        pos = _f64(pos_lid).reshape(3)
        boxes = np.zeros((3, 6), np.float32)
        nb = self.L.orc_fov_segment(self.h, _p(pos, _f64p), _p(boxes, _f32p))
        return boxes[:nb]


def esti_plane(pts5x3, threshold=0.1):
    p = _f32(pts5x3).reshape(5, 3)
    out = np.zeros(4, np.float32)
    ok = lib().orc_esti_plane(_p(p, _f32p), threshold, _p(out, _f32p))
    return bool(ok), out


def boxplus(state, dx):
    s = _f64(state).reshape(36).copy()
    d = _f64(dx).reshape(23)
    lib().orc_boxplus(_p(s, _f64p), _p(d, _f64p))
    return s


def imu_predict(state, P, dt, acc, gyro, Q):
    """esekfom::predict restatement (orc_imu.cpp); returns (state, P)."""
    s = _f64(state).reshape(36).copy()
    P = _f64(P).reshape(23, 23).copy()
    lib().orc_imu_predict(_p(s, _f64p), _p(P, _f64p), float(dt), _p(_f64(acc).reshape(3), _f64p), _p(_f64(gyro).reshape(3), _f64p),
                          _p(_f64(Q).reshape(12, 12), _f64p))
    return s, P


def undistort(xyz, t_ms, imu_pose, end_state):
    """UndistortPcl's backward sweep (orc_imu.cpp) on a time-sorted cloud; returns the compensated copy."""
    out = _f32(xyz).reshape(-1, 3).copy()
    t = _f32(t_ms).reshape(-1)
    pose = _f64(imu_pose).reshape(-1, 22)
    lib().orc_undistort(_p(out, _f32p), _p(t, _f32p), len(out), _p(pose, _f64p), len(pose), _p(_f64(end_state).reshape(36), _f64p))
    return out


def boxminus(a, b):
    a = _f64(a).reshape(36)
    b = _f64(b).reshape(36)
    d = np.zeros(23)
    lib().orc_boxminus(_p(a, _f64p), _p(b, _f64p), _p(d, _f64p))
    return d


class OracleLoop:
    """CPU restatement of LoopClosure (loop_closure.cpp) with PCL's ICP / VoxelGrid semantics."""

    def __init__(self, max_corr_dist=52.5, trans_eps=0.01, fit_eps=0.01, max_iter=50, nthreads=1):
        self.L = lib()
        self.h = C.c_void_p(self.L.orc_loop_create(max_corr_dist, trans_eps, fit_eps, max_iter, nthreads))

    def close(self):
        if self.h:
            self.L.orc_loop_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def add_keyframe(self, xyzi, pose, stamp=0.0):
        xyzi = _f32(xyzi).reshape(-1, 4)
        pose = _f64(pose).reshape(16)
        return self.L.orc_loop_add_keyframe(self.h, _p(xyzi, _f32p), len(xyzi), _p(pose, _f64p), stamp)

    def set_pose(self, idx, pose):
        pose = _f64(pose).reshape(16)
        self.L.orc_loop_set_pose(self.h, int(idx), _p(pose, _f64p))

    def fetch_closest(self, query_idx, radius, tdiff):
        return self.L.orc_loop_fetch_closest(self.h, query_idx, radius, tdiff)

    def set_src_dst(self, src_idx, dst_idx, rng, leaf, cap=1 << 22):
        src = np.zeros((cap, 4), np.float32)
        dst = np.zeros((cap, 4), np.float32)
        ns, nd = C.c_int(0), C.c_int(0)
        self.L.orc_loop_set_src_dst(self.h, src_idx, dst_idx, rng, leaf, _p(src, _f32p), C.byref(ns), _p(dst, _f32p), C.byref(nd), cap)
        return src[: ns.value].copy(), dst[: nd.value].copy()

    def icp_align(self, src, dst, want_aligned=False):
        src = _f32(src).reshape(-1, 4)
        dst = _f32(dst).reshape(-1, 4)
        T = np.zeros((4, 4), np.float32)
        score = C.c_double(0)
        conv, it, st = C.c_int(0), C.c_int(0), C.c_int(0)
        aligned = np.zeros_like(src) if want_aligned else None
        corr = np.zeros(len(src), np.int32) if want_aligned else None
        self.L.orc_icp_align(self.h, _p(src, _f32p), len(src), _p(dst, _f32p), len(dst), _p(T, _f32p), C.byref(score), C.byref(conv),
                             C.byref(it), C.byref(st), _p(aligned, _f32p), _p(corr, _i32p))
        out = dict(T=T, score=score.value, converged=bool(conv.value), iterations=it.value, state=st.value)
        if want_aligned:
            out["aligned"] = aligned
            out["corr"] = corr
        return out

    def verify(self, src_idx, dst_idx, rng, leaf, score_thr):
        T = np.zeros((4, 4), np.float32)
        score = C.c_double(0)
        conv, valid, it, ns, nd = (C.c_int(0) for _ in range(5))
        self.L.orc_loop_verify(self.h, src_idx, dst_idx, rng, leaf, score_thr, _p(T, _f32p), C.byref(score), C.byref(conv),
                               C.byref(valid), C.byref(it), C.byref(ns), C.byref(nd))
        return dict(T=T, score=score.value, converged=bool(conv.value), valid=bool(valid.value), iterations=it.value,
                    n_src=ns.value, n_dst=nd.value)


def transform_pcd(xyzi, T):
    xyzi = _f32(xyzi).reshape(-1, 4)
    T = _f64(T).reshape(16)
    out = np.zeros_like(xyzi)
    lib().orc_transform_pcd(_p(xyzi, _f32p), len(xyzi), _p(T, _f64p), _p(out, _f32p))
    return out


def pose_inverse(T):
    """pose_eig_.inverse() of the PosePcd constructor (pose_pcd.hpp:38)."""
    T = _f64(T).reshape(16)
    out = np.zeros(16)
    if not lib().orc_pose_inverse(_p(T, _f64p), _p(out, _f64p)):
        raise ValueError("singular pose")
    return out.reshape(4, 4)


def keyframe_from_world(xyzi_world, pose):
    """PosePcd constructor: pcd_ = transformPcd(cloud_world, pose.inverse()) (pose_pcd.hpp:37-39)."""
    return transform_pcd(xyzi_world, pose_inverse(pose))


def voxelize(xyzi, leaf):
    xyzi = _f32(xyzi).reshape(-1, 4)
    out = np.zeros_like(xyzi)
    n = lib().orc_voxelize(_p(xyzi, _f32p), len(xyzi), leaf, _p(out, _f32p))
    return out[:n].copy()


def svd3(M):
    M = _f32(M).reshape(9)
    U = np.zeros(9, np.float32)
    S = np.zeros(3, np.float32)
    V = np.zeros(9, np.float32)
    lib().orc_svd3(_p(M, _f32p), _p(U, _f32p), _p(S, _f32p), _p(V, _f32p))
    return U.reshape(3, 3), S, V.reshape(3, 3)
