"""B200-native FAST-LIO measurement path + loop-closure verification: Python bindings of the C-ABI.

The product is the CUDA shared library ``lib/liblio_b200.so`` (sources in ``csrc/``, interface in
``include/lio_b200.h`` / ``include/loop_b200.h``).  This package is only the ctypes view of it that
the tests and ``bench.py`` use; the reference-facing host code is C++ (``adapters/``).

There is no CPU fallback: importing works without a GPU (so symbol checks can run), but every
compute entry point raises :class:`LioError` when the library or a B200 is missing.

The directory name contains '-' and '.', so import it with
``importlib.import_module("fast-lio-sam_gps_b200")``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "liblio_b200.so")
CSRC_DIR = os.path.join(_HERE, "csrc")

LIO_OK = 0
STATE_DOUBLES = 36
POSE6D_DOUBLES = 22
ERR_DOF = 23
NUM_MATCH = 5


class LioError(RuntimeError):
    pass


class LioConfig(C.Structure):
    _fields_ = [
        ("device", C.c_int),
        ("filter_size_map", C.c_float),
        ("cell_mult", C.c_int),
        ("max_points", C.c_uint32),
        ("max_cells", C.c_uint32),
        ("max_scan", C.c_uint32),
        ("max_sqdist", C.c_float),
        ("extrinsic_est_en", C.c_int),
        ("cube_side_length", C.c_double),
        ("det_range", C.c_float),
        ("stream", C.c_void_p),
    ]


def build_library(verbose: bool = False) -> str:
    """Compile csrc/ for sm_100a with nvcc (in-tree, so the .so travels with the repo snapshot)."""
    res = subprocess.run(["make", "-C", CSRC_DIR], capture_output=True, text=True)
    if res.returncode != 0:
        raise LioError("nvcc build failed:\n" + res.stdout[-4000:] + res.stderr[-4000:])
    if verbose:
        print(res.stdout[-2000:])
    return LIB_PATH


_lib = None

_f32p = C.POINTER(C.c_float)
_f64p = C.POINTER(C.c_double)
_i32p = C.POINTER(C.c_int)
_u32p = C.POINTER(C.c_uint32)
_u8p = C.POINTER(C.c_uint8)

# name -> (restype, argtypes); every symbol include/lio_b200.h declares
LIO_SYMBOLS = {
    "lio_version": (C.c_char_p, []),
    "lio_last_error": (C.c_char_p, []),
    "lio_device_count": (C.c_int, []),
    "lio_ctx_create": (C.c_int, [C.POINTER(LioConfig), C.POINTER(C.c_void_p)]),
    "lio_ctx_destroy": (None, [C.c_void_p]),
    "lio_ctx_sync": (C.c_int, [C.c_void_p]),
    "lio_map_build": (C.c_int, [C.c_void_p, _f32p, C.c_int]),
    "lio_map_add": (C.c_int, [C.c_void_p, _f32p, C.c_int, C.c_int, _i32p]),
    "lio_map_delete_boxes": (C.c_int, [C.c_void_p, _f32p, C.c_int, _i32p]),
    "lio_map_size": (C.c_int, [C.c_void_p, _i32p]),
    "lio_map_export": (C.c_int, [C.c_void_p, _f32p, _u32p, C.c_int, _i32p]),
    "lio_map_knn": (C.c_int, [C.c_void_p, _f32p, C.c_int, C.c_int, C.c_float, _u32p, _f32p, _f32p, _i32p]),
    "lio_scan_upload": (C.c_int, [C.c_void_p, _f32p, C.c_int]),
    "lio_scan_prefetch": (C.c_int, [C.c_void_p, _f32p, C.c_int]),
    "lio_scan_set_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "lio_h_share": (C.c_int, [C.c_void_p, _f64p, C.c_int, _f64p, _f64p, _i32p, _f64p]),
    "lio_h_share_rows": (C.c_int, [C.c_void_p, _f64p, C.c_int, _u8p, _f64p, _f64p, _i32p]),
    "lio_scan_get": (C.c_int, [C.c_void_p, _f32p, _u32p, _f32p, _i32p, _u8p, _f32p]),
    "lio_iekf_update": (C.c_int, [C.c_void_p, _f64p, _f64p, C.c_double, C.c_int, _f64p, _f64p, _i32p, _i32p, _i32p]),
    "lio_iekf_update_map": (C.c_int, [C.c_void_p, _f64p, _f64p, C.c_double, C.c_int, _f64p, C.c_int, _f64p, _i32p, _i32p, _i32p]),
    "lio_imu_predict": (C.c_int, [_f64p, _f64p, C.c_double, _f64p, _f64p, _f64p]),
    "lio_scan_undistort": (C.c_int, [C.c_void_p, _f32p, _f64p, C.c_int, _f64p]),
    "lio_scan_get_body": (C.c_int, [C.c_void_p, _f32p]),
    "lio_map_incremental": (C.c_int, [C.c_void_p, _f64p, C.c_int, _i32p, _i32p]),
    "lio_fov_segment": (C.c_int, [C.c_void_p, _f64p, _f32p, _i32p]),
    "lio_boxplus": (None, [_f64p, _f64p]),
    "lio_boxminus": (None, [_f64p, _f64p, _f64p]),
    "lio_timing_enable": (C.c_int, [C.c_void_p, C.c_int]),
    "lio_timing_get": (C.c_int, [C.c_void_p, C.c_char_p, _f64p, _i32p]),
    "lio_timing_reset": (C.c_int, [C.c_void_p]),
    "lio_knn_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "lio_knn_stats_detail": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "lio_map_compact": (C.c_int, [C.c_void_p, C.c_int]),
    "lio_map_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "lio_tile_profile": (C.c_int, [C.c_void_p, _u32p, C.c_int, _i32p]),
    "lio_gather_probe": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "lio_launch_count": (C.c_int64, [C.c_void_p]),
    "lio_hth_reduce_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
}


def load_library():
    """dlopen the in-tree library and bind every declared symbol.  Raises LioError if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LioError(f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in LIO_SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    try:
        from . import loop as _loop  # noqa: WPS433  (binds include/loop_b200.h when present)

        _loop.bind(lib)
    except ImportError:
        pass
    _lib = lib
    return lib


def _p(a, ty):
    return None if a is None else a.ctypes.data_as(ty)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class LioContext:
    """GPU map + current scan (mirror of the file-scope globals h_share_model uses)."""

    def __init__(self, filter_size_map=0.5, max_points=1 << 20, max_scan=1 << 17, max_cells=0, cell_mult=2, max_sqdist=5.0,
                 extrinsic_est_en=False, cube_side_length=1000.0, det_range=300.0, device=0, stream=None):
        self.lib = load_library()
        self.cfg = LioConfig(device, filter_size_map, cell_mult, max_points, max_cells, max_scan, max_sqdist, int(extrinsic_est_en),
                             cube_side_length, det_range, stream)
        self.h = C.c_void_p()
        self._check(self.lib.lio_ctx_create(C.byref(self.cfg), C.byref(self.h)))
        self.scan_n = 0

    def _check(self, rc):
        if rc != LIO_OK:
            raise LioError(f"lio error {rc}: {self.lib.lio_last_error().decode()}")

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.lib.lio_ctx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- map
    def map_build(self, xyz):
        xyz = _f32(xyz).reshape(-1, 3)
        self._check(self.lib.lio_map_build(self.h, _p(xyz, _f32p), len(xyz)))

    def map_add(self, xyz, downsample=True):
        xyz = _f32(xyz).reshape(-1, 3)
        net = C.c_int(0)
        self._check(self.lib.lio_map_add(self.h, _p(xyz, _f32p), len(xyz), int(downsample), C.byref(net)))
        return net.value

    def map_delete_boxes(self, boxes):
        boxes = _f32(boxes).reshape(-1, 6)
        n = C.c_int(0)
        self._check(self.lib.lio_map_delete_boxes(self.h, _p(boxes, _f32p), len(boxes), C.byref(n)))
        return n.value

    def map_size(self):
        n = C.c_int(0)
        self._check(self.lib.lio_map_size(self.h, C.byref(n)))
        return n.value

    def map_export(self):
        n = self.map_size()
        xyz = np.zeros((max(n, 1), 3), np.float32)
        ids = np.zeros(max(n, 1), np.uint32)
        tot = C.c_int(0)
        self._check(self.lib.lio_map_export(self.h, _p(xyz, _f32p), _p(ids, _u32p), n, C.byref(tot)))
        return xyz[: tot.value], ids[: tot.value]

    def map_knn(self, q, k=5, max_sqdist=5.0):
        q = _f32(q).reshape(-1, 3)
        n = len(q)
        ids = np.zeros((n, k), np.uint32)
        d2 = np.zeros((n, k), np.float32)
        pts = np.zeros((n, k, 3), np.float32)
        cnt = np.zeros(n, np.int32)
        self._check(self.lib.lio_map_knn(self.h, _p(q, _f32p), n, k, max_sqdist, _p(ids, _u32p), _p(d2, _f32p), _p(pts, _f32p),
                                         _p(cnt, _i32p)))
        return ids, d2, pts, cnt

    # ---- scan + measurement model
    def scan_upload(self, body_xyz):
        body_xyz = _f32(body_xyz).reshape(-1, 3)
        self._keep = body_xyz  # the copy is asynchronous
        self.scan_n = len(body_xyz)
        self._check(self.lib.lio_scan_upload(self.h, _p(body_xyz, _f32p), len(body_xyz)))

    def scan_upload_ptr(self, host_ptr, n):
        """Upload from a raw (ideally pinned) host pointer to n*3 floats."""
        self.scan_n = n
        self._check(self.lib.lio_scan_upload(self.h, C.cast(host_ptr, _f32p), n))

    def scan_prefetch_ptr(self, host_ptr, n):
        """Start the H2D copy of the next scan (pinned host pointer) on the copy stream; see lio_scan_prefetch."""
        self._check(self.lib.lio_scan_prefetch(self.h, C.cast(host_ptr, _f32p), n))

    def scan_set_device(self, dev_ptr, n):
        self.scan_n = n
        self._check(self.lib.lio_scan_set_device(self.h, C.c_void_p(dev_ptr), n))

    def h_share(self, state, redo_knn=True):
        state = _f64(state).reshape(STATE_DOUBLES)
        HTH = np.zeros((12, 12))
        HTz = np.zeros(12)
        n_eff = C.c_int(0)
        res = C.c_double(0)
        self._check(self.lib.lio_h_share(self.h, _p(state, _f64p), int(redo_knn), _p(HTH, _f64p), _p(HTz, _f64p), C.byref(n_eff),
                                         C.byref(res)))
        return HTH, HTz, n_eff.value, res.value

    def h_share_rows(self, state, redo_knn=True):
        state = _f64(state).reshape(STATE_DOUBLES)
        n = self.scan_n
        sel = np.zeros(n, np.uint8)
        rows = np.zeros((n, 12))
        z = np.zeros(n)
        n_eff = C.c_int(0)
        self._check(self.lib.lio_h_share_rows(self.h, _p(state, _f64p), int(redo_knn), _p(sel, _u8p), _p(rows, _f64p), _p(z, _f64p),
                                              C.byref(n_eff)))
        return sel, rows, z, n_eff.value

    def scan_get(self):
        n = self.scan_n
        out = dict(world=np.zeros((n, 3), np.float32), nn_ids=np.zeros((n, 5), np.uint32), nn_xyz=np.zeros((n, 5, 3), np.float32),
                   nn_cnt=np.zeros(n, np.int32), sel=np.zeros(n, np.uint8), plane=np.zeros((n, 4), np.float32))
        self._check(self.lib.lio_scan_get(self.h, _p(out["world"], _f32p), _p(out["nn_ids"], _u32p), _p(out["nn_xyz"], _f32p),
                                          _p(out["nn_cnt"], _i32p), _p(out["sel"], _u8p), _p(out["plane"], _f32p)))
        return out

    def iekf_update(self, state, P, R=0.001, max_iter=3, limit=None, trace=False, with_map=None):
        """esekfom's iterated update; with_map=True/False also runs map_incremental(flg_ekf_inited=with_map... see
        iekf_update_map) in the same enqueue (lio_iekf_update_map)."""
        state = _f64(state).reshape(STATE_DOUBLES).copy()
        P = _f64(P).reshape(ERR_DOF, ERR_DOF).copy()
        limit = _f64(np.full(ERR_DOF, 0.001) if limit is None else limit)
        it_states = np.zeros((max_iter + 1, STATE_DOUBLES))
        n_eff = np.zeros(max_iter + 1, np.int32)
        knn = np.zeros(max_iter + 1, np.int32)
        done = C.c_int(0)
        tr = (_p(it_states, _f64p) if trace else None, _p(n_eff, _i32p) if trace else None, _p(knn, _i32p) if trace else None, C.byref(done))
        if with_map is None:
            self._check(self.lib.lio_iekf_update(self.h, _p(state, _f64p), _p(P, _f64p), R, max_iter, _p(limit, _f64p), *tr))
        else:
            self._check(self.lib.lio_iekf_update_map(self.h, _p(state, _f64p), _p(P, _f64p), R, max_iter, _p(limit, _f64p),
                                                     int(with_map), *tr))
        if trace:
            d = done.value
            return state, P, dict(iters=d, states=it_states[:d], n_eff=n_eff[:d], knn=knn[:d])
        return state, P, done.value

    def scan_undistort(self, t_ms, imu_pose, end_state):
        """UndistortPcl on the uploaded scan, in place (t_ms: per-point curvature; imu_pose: (n_pose, 22) Pose6D rows)."""
        t_ms = _f32(t_ms).reshape(-1)
        pose = _f64(imu_pose).reshape(-1, POSE6D_DOUBLES)
        assert len(t_ms) == self.scan_n
        self._check(self.lib.lio_scan_undistort(self.h, _p(t_ms, _f32p), _p(pose, _f64p), len(pose), _p(_f64(end_state).reshape(STATE_DOUBLES), _f64p)))

    def scan_body(self):
        out = np.zeros((self.scan_n, 3), np.float32)
        self._check(self.lib.lio_scan_get_body(self.h, _p(out, _f32p)))
        return out

    def iekf_update_map(self, state, P, flg_ekf_inited=True, **kw):
        """The per-scan body of the reference's main loop: iterated update, then map_incremental() at the updated state."""
        return self.iekf_update(state, P, with_map=bool(flg_ekf_inited), **kw)

    def map_incremental(self, state, flg_ekf_inited=True, want_counts=True):
        state = _f64(state).reshape(STATE_DOUBLES)
        a, b = C.c_int(0), C.c_int(0)
        self._check(self.lib.lio_map_incremental(self.h, _p(state, _f64p), int(flg_ekf_inited), C.byref(a) if want_counts else None,
                                                 C.byref(b) if want_counts else None))
        return a.value, b.value

    def fov_segment(self, pos_lid):
        pos = _f64(pos_lid).reshape(3)
        boxes = np.zeros((3, 6), np.float32)
        nb = C.c_int(0)
        self._check(self.lib.lio_fov_segment(self.h, _p(pos, _f64p), _p(boxes, _f32p), C.byref(nb)))
        return boxes[: nb.value]

    def sync(self):
        self._check(self.lib.lio_ctx_sync(self.h))

    # ---- measurement hooks
    def timing(self, on=True):
        self.lib.lio_timing_enable(self.h, int(on))

    def timing_reset(self):
        self.lib.lio_timing_reset(self.h)

    def timing_get(self, name):
        ms, n = C.c_double(0), C.c_int(0)
        self._check(self.lib.lio_timing_get(self.h, name.encode(), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def knn_fallback_queries(self):
        v = C.c_uint64(0)
        self._check(self.lib.lio_knn_stats(self.h, C.byref(v)))
        return int(v.value)

    def knn_stats_detail(self):
        v = (C.c_uint64 * 4)()
        self._check(self.lib.lio_knn_stats_detail(self.h, v))
        return dict(per_query=int(v[0]), overflow=int(v[1]), incoherent=int(v[2]))

    def map_compact(self, renumber=False):
        """Rebuild the store from its live points (reclaims cells / buckets emptied by box deletes)."""
        self._check(self.lib.lio_map_compact(self.h, 1 if renumber else 0))

    def map_stats(self):
        v = (C.c_uint64 * 8)()
        self._check(self.lib.lio_map_stats(self.h, v))
        return dict(live=int(v[0]), next_id=int(v[1]), pool_top=int(v[2]), pool_cap=int(v[3]), n_cells=int(v[4]), cell_limit=int(v[5]),
                    error=int(v[6]), dead_since_compact=int(v[7]))

    def tile_profile(self):
        """(tiles, 8) uint32 per-tile counters of the last scan search (needs LIO_TILE_PROF=1 at context creation)."""
        n = C.c_int(0)
        self._check(self.lib.lio_tile_profile(self.h, None, 0, C.byref(n)))
        out = np.zeros((max(n.value, 1), 8), np.uint32)
        self._check(self.lib.lio_tile_profile(self.h, _p(out, _u32p), n.value, C.byref(n)))
        return out[: n.value]

    def gather_probe(self, table_bytes=320 << 20, reps=5):
        a, b = C.c_double(0), C.c_double(0)
        self._check(self.lib.lio_gather_probe(self.h, table_bytes, reps, C.byref(a), C.byref(b)))
        return a.value, b.value

    def launch_count(self):
        return int(self.lib.lio_launch_count(self.h))


STATIC_REDUCE_MAX_ROWS = 1 << 20  # LIO_STATIC_REDUCE_MAX_ROWS


def hth_reduce_static(rows_dev, z_dev, n, hth_dev, htz_dev, stream=None):
    """lio_hth_reduce_static: rows^T rows / rows^T z from DEVICE addresses (ints), with no allocation on the way; the
    per-block sums live in the library's static device array (shared by every call on the GPU).  Asynchronous on `stream`."""
    lib = load_library()
    rc = lib.lio_hth_reduce_static(rows_dev, z_dev, int(n), hth_dev, htz_dev, stream)
    if rc != LIO_OK:
        raise LioError(f"lio_hth_reduce_static failed ({rc}): {lib.lio_last_error().decode()}")


def imu_predict(state, P, dt, acc, gyro, Q):
    """esekfom::predict on the host (lio_imu_predict); returns (state, P)."""
    lib = load_library()
    state = _f64(state).reshape(STATE_DOUBLES).copy()
    P = _f64(P).reshape(ERR_DOF, ERR_DOF).copy()
    rc = lib.lio_imu_predict(_p(state, _f64p), _p(P, _f64p), float(dt), _p(_f64(acc).reshape(3), _f64p), _p(_f64(gyro).reshape(3), _f64p),
                             _p(_f64(Q).reshape(12, 12), _f64p))
    if rc != LIO_OK:
        raise LioError(f"lio_imu_predict: {rc}")
    return state, P


def identity_state(pos=(0, 0, 0), rot=None, grav=(0.19614, -0.14711, -9.80593)):
    """LIO_STATE_DOUBLES vector: pos | rot | offset_R_L_I | offset_T_L_I | vel | bg | ba | grav."""
    s = np.zeros(STATE_DOUBLES)
    s[0:3] = pos
    s[3:12] = (np.eye(3) if rot is None else np.asarray(rot, float)).reshape(9)
    s[12:21] = np.eye(3).reshape(9)
    s[33:36] = grav
    return s
