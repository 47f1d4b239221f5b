"""ctypes view of include/loop_b200.h — the B200 loop-closure verification path — and a host-side mirror of the
reference's `LoopClosure` class (fast_lio_sam/include/loop_closure.h:39-72) with the same method names and error
behaviour (never raises for a rejected match: `is_valid=False`, `score` still set), used by tests/ and bench.py.

No CPU fallback: constructing a `LoopClosure` raises LioError when the CUDA library or a B200 is missing.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

_f32p = C.POINTER(C.c_float)
_f64p = C.POINTER(C.c_double)
_i32p = C.POINTER(C.c_int)
_i64p = C.POINTER(C.c_int64)


class LoopConfig(C.Structure):
    _fields_ = [
        ("device", C.c_int),
        ("max_corr_dist", C.c_double),
        ("transformation_epsilon", C.c_double),
        ("euclidean_fitness_epsilon", C.c_double),
        ("max_iterations", C.c_int),
        ("rotation_epsilon", C.c_double),
        ("mse_threshold_absolute", C.c_double),
        ("max_similar_iterations", C.c_int),
        ("min_correspondences", C.c_int),
        ("stream", C.c_void_p),
    ]


class LoopResult(C.Structure):
    _fields_ = [
        ("T", C.c_float * 16),
        ("score", C.c_double),
        ("converged", C.c_int32),
        ("valid", C.c_int32),
        ("iterations", C.c_int32),
        ("state", C.c_int32),
        ("n_src", C.c_int32),
        ("n_dst", C.c_int32),
    ]


class LoopPair(C.Structure):
    _fields_ = [("src_idx", C.c_int32), ("dst_idx", C.c_int32)]


_resp = C.POINTER(LoopResult)

# name -> (restype, argtypes); every symbol include/loop_b200.h declares
LOOP_SYMBOLS = {
    "loop_ctx_create": (C.c_int, [C.POINTER(LoopConfig), C.POINTER(C.c_void_p)]),
    "loop_ctx_destroy": (None, [C.c_void_p]),
    "loop_last_error": (C.c_char_p, []),
    "loop_keyframe_add": (C.c_int, [C.c_void_p, _f32p, C.c_int, _f64p, C.c_double, _i32p]),
    "loop_keyframe_set_pose": (C.c_int, [C.c_void_p, C.c_int, _f64p]),
    "loop_keyframe_add_world": (C.c_int, [C.c_void_p, _f32p, C.c_int, _f64p, C.c_double, _i32p]),
    "loop_keyframe_set_poses": (C.c_int, [C.c_void_p, _i32p, _f64p, C.c_int]),
    "loop_fetch_closest_batch": (C.c_int, [C.c_void_p, _i32p, C.c_int, C.c_double, C.c_double, _i32p]),
    "loop_keyframe_count": (C.c_int, [C.c_void_p]),
    "loop_keyframe_clear": (C.c_int, [C.c_void_p]),
    "loop_fetch_closest": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_double, _i32p]),
    "loop_transform_pcd": (C.c_int, [C.c_void_p, _f32p, C.c_int, _f64p, _f32p]),
    "loop_voxelize": (C.c_int, [C.c_void_p, _f32p, C.c_int, C.c_float, _f32p, _i32p]),
    "loop_set_src_dst": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, _f32p, _i32p, _f32p, _i32p, C.c_int]),
    "loop_icp_align": (C.c_int, [C.c_void_p, _f32p, C.c_int, _f32p, C.c_int, C.c_double, _resp, _f32p, _i32p]),
    "loop_verify_batch": (C.c_int, [C.c_void_p, C.POINTER(LoopPair), C.c_int, C.c_int, C.c_float, C.c_double, _resp]),
    "loop_get_cloud": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _f32p, C.c_int, _i32p]),
    "loop_timing_enable": (C.c_int, [C.c_void_p, C.c_int]),
    "loop_timing_get": (C.c_int, [C.c_void_p, C.c_char_p, _f64p, _i32p]),
    "loop_timing_reset": (C.c_int, [C.c_void_p]),
    "loop_launch_count": (C.c_int64, [C.c_void_p]),
    "loop_batch_stats": (C.c_int, [C.c_void_p, _i64p, _i64p, _i64p]),
}


def bind(lib):
    for name, (res, args) in LOOP_SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def _p(a, ty):
    return None if a is None else a.ctypes.data_as(ty)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _result(r: LoopResult):
    return dict(T=np.array(r.T, np.float32).reshape(4, 4), score=r.score, converged=bool(r.converged), valid=bool(r.valid),
                iterations=r.iterations, state=r.state, n_src=r.n_src, n_dst=r.n_dst)


def shard_pairs(n_pairs, rank, world):
    """Contiguous slice [lo, hi) of the pair list owned by `rank` (sizes differ by at most one)."""
    base, rem = divmod(n_pairs, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_pair_indices(n_pairs, rank, world):
    """Pair indices owned by `rank`: the list is cut into 2*world contiguous blocks and rank r takes blocks r and
    2*world-1-r (dealt there and back).  Contiguous blocks keep the pairs of neighbouring query keyframes together (they
    share submaps, so each is built once per rank); the mirrored second block cancels a trend in ICP iteration count along
    the list (drift grows along a drive), which with one contiguous slice per rank left the slowest rank 25 % over the mean."""
    out = []
    for b in (rank, 2 * world - 1 - rank):
        lo, hi = shard_pairs(n_pairs, b, 2 * world)
        out.extend(range(lo, hi))
    return out


class LoopClosure:
    """Mirror of the reference's LoopClosure: keyframes live on the device, candidate pairs are verified in batches.

    Constructor arguments are LoopClosureConfig's fields (loop_closure.h:21-29) and the ICP constants of
    loop_closure.cpp:7-11; the defaults are the effective values of config.yaml / fast_lio_sam.cpp:64-76.
    """

    def __init__(self, num_submap_keyframes=5, voxel_res=0.3, loop_detection_radius=35.0, loop_detection_timediff_threshold=30.0,
                 icp_score_threshold=1.5, icp_max_corr_dist=None, transformation_epsilon=0.01, euclidean_fitness_epsilon=0.01,
                 max_iterations=50, device=0, stream=None):
        from . import LioError, load_library

        self._err = LioError
        self.lib = load_library()
        self.num_submap_keyframes = num_submap_keyframes
        self.voxel_res = voxel_res
        self.loop_detection_radius = loop_detection_radius
        self.loop_detection_timediff_threshold = loop_detection_timediff_threshold
        self.icp_score_threshold = icp_score_threshold
        self.icp_max_corr_dist = loop_detection_radius * 1.5 if icp_max_corr_dist is None else icp_max_corr_dist
        cfg = LoopConfig(device, self.icp_max_corr_dist, transformation_epsilon, euclidean_fitness_epsilon, max_iterations, 0.0, 0.0, 0, 0,
                         stream)
        self.h = C.c_void_p()
        self._check(self.lib.loop_ctx_create(C.byref(cfg), C.byref(self.h)))

    def _check(self, rc):
        if rc != 0:
            raise self._err(f"loop error {rc}: {self.lib.loop_last_error().decode()}")

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.lib.loop_ctx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- keyframes_ ---------------------------------------------------------------------------------------------
    def add_keyframe(self, xyzi, pose_corrected, stamp=0.0):
        xyzi = _f32(xyzi).reshape(-1, 4)
        pose = _f64(pose_corrected).reshape(16)
        idx = C.c_int(-1)
        self._check(self.lib.loop_keyframe_add(self.h, _p(xyzi, _f32p), len(xyzi), _p(pose, _f64p), float(stamp), C.byref(idx)))
        return idx.value

    def add_keyframe_world(self, xyzi_world, pose, stamp=0.0):
        """PosePcd's constructor (pose_pcd.hpp:25-43): world-frame cloud in, LiDAR-frame cloud stored (pose^-1 on the device)."""
        xyzi = _f32(xyzi_world).reshape(-1, 4)
        pose = _f64(pose).reshape(16)
        idx = C.c_int(-1)
        self._check(self.lib.loop_keyframe_add_world(self.h, _p(xyzi, _f32p), len(xyzi), _p(pose, _f64p), float(stamp), C.byref(idx)))
        return idx.value

    def set_poses(self, idx, poses_corrected):
        """pose_corrected_eig_ of several keyframes at once (after a pose-graph update)."""
        idx = np.ascontiguousarray(idx, np.int32).reshape(-1)
        poses = _f64(poses_corrected).reshape(-1, 16)
        assert len(idx) == len(poses)
        self._check(self.lib.loop_keyframe_set_poses(self.h, _p(idx, _i32p), _p(poses, _f64p), len(idx)))

    def fetch_closest_keyframe_idx_batch(self, query_idx):
        """fetchClosestKeyframeIdx for many query keyframes in one device launch."""
        q = np.ascontiguousarray(query_idx, np.int32).reshape(-1)
        out = np.full(len(q), -1, np.int32)
        self._check(self.lib.loop_fetch_closest_batch(self.h, _p(q, _i32p), len(q), self.loop_detection_radius,
                                                      self.loop_detection_timediff_threshold, _p(out, _i32p)))
        return out

    def set_pose(self, idx, pose_corrected):
        pose = _f64(pose_corrected).reshape(16)
        self._check(self.lib.loop_keyframe_set_pose(self.h, idx, _p(pose, _f64p)))

    def __len__(self):
        return self.lib.loop_keyframe_count(self.h)

    def clear(self):
        self._check(self.lib.loop_keyframe_clear(self.h))

    # ---- LoopClosure's methods ------------------------------------------------------------------------------------
    def fetch_closest_keyframe_idx(self, query_idx):
        out = C.c_int(-1)
        self._check(self.lib.loop_fetch_closest(self.h, query_idx, self.loop_detection_radius, self.loop_detection_timediff_threshold,
                                                C.byref(out)))
        return out.value

    def set_src_and_dst_cloud(self, src_idx, dst_idx, submap_range=None, voxel_res=None, cap=1 << 22):
        rng = self.num_submap_keyframes if submap_range is None else submap_range
        leaf = self.voxel_res if voxel_res is None else voxel_res
        ns, nd = C.c_int(0), C.c_int(0)
        self._check(self.lib.loop_set_src_dst(self.h, src_idx, dst_idx, rng, leaf, None, C.byref(ns), None, C.byref(nd), 0))
        src = np.zeros((max(ns.value, 1), 4), np.float32)
        dst = np.zeros((max(nd.value, 1), 4), np.float32)
        cap = max(ns.value, nd.value, 1)
        self._check(self.lib.loop_set_src_dst(self.h, src_idx, dst_idx, rng, leaf, _p(src, _f32p), C.byref(ns), _p(dst, _f32p), C.byref(nd),
                                              cap))
        return src[: ns.value], dst[: nd.value]

    def icp_alignment(self, src, dst, want_aligned=False):
        src = _f32(src).reshape(-1, 4)
        dst = _f32(dst).reshape(-1, 4)
        r = LoopResult()
        aligned = np.zeros_like(src) if want_aligned else None
        corr = np.zeros(len(src), np.int32) if want_aligned else None
        self._check(self.lib.loop_icp_align(self.h, _p(src, _f32p), len(src), _p(dst, _f32p), len(dst), self.icp_score_threshold, C.byref(r),
                                            _p(aligned, _f32p), _p(corr, _i32p)))
        out = _result(r)
        if want_aligned:
            out["aligned"] = aligned
            out["corr"] = corr
        return out

    def perform_loop_closure(self, query_idx, closest_idx=None):
        """performLoopClosure(query, keyframes[, closest]) for ONE candidate (loop_closure.cpp:94-126)."""
        if closest_idx is None:
            closest_idx = self.fetch_closest_keyframe_idx(query_idx)
        if closest_idx < 0:
            return dict(T=np.eye(4, dtype=np.float32), score=float(np.finfo(np.float64).max), converged=False, valid=False, iterations=0,
                        state=0, n_src=0, n_dst=0)
        return self.verify_batch([(query_idx, closest_idx)])[0]

    def verify_batch(self, pairs, submap_range=None, voxel_res=None):
        """performLoopClosure for many (query, closest) pairs in one device batch."""
        rng = self.num_submap_keyframes if submap_range is None else submap_range
        leaf = self.voxel_res if voxel_res is None else voxel_res
        n = len(pairs)
        arr = (LoopPair * max(n, 1))()
        for i, (a, b) in enumerate(pairs):
            arr[i].src_idx, arr[i].dst_idx = int(a), int(b)
        res = (LoopResult * max(n, 1))()
        self._check(self.lib.loop_verify_batch(self.h, arr, n, rng, leaf, self.icp_score_threshold, res))
        return [_result(res[i]) for i in range(n)]

    def verify_batch_raw(self, pair_array, n, res_array, submap_range=None, voxel_res=None):
        """Same call with caller-owned ctypes arrays (bench.py: no per-call Python marshalling)."""
        rng = self.num_submap_keyframes if submap_range is None else submap_range
        leaf = self.voxel_res if voxel_res is None else voxel_res
        self._check(self.lib.loop_verify_batch(self.h, pair_array, n, rng, leaf, self.icp_score_threshold, res_array))

    def get_cloud(self, pair, which):
        """which: 'src' | 'dst' | 'aligned' — the debug getters of loop_closure.h:66-68 for a pair of the last batch."""
        w = {"src": 0, "dst": 1, "aligned": 2}[which]
        n = C.c_int(0)
        self._check(self.lib.loop_get_cloud(self.h, pair, w, None, 0, C.byref(n)))
        out = np.zeros((max(n.value, 1), 4), np.float32)
        self._check(self.lib.loop_get_cloud(self.h, pair, w, _p(out, _f32p), n.value, C.byref(n)))
        return out[: n.value]

    # ---- utilities.hpp ------------------------------------------------------------------------------------------------
    def transform_pcd(self, xyzi, T):
        xyzi = _f32(xyzi).reshape(-1, 4)
        T = _f64(T).reshape(16)
        out = np.zeros_like(xyzi)
        self._check(self.lib.loop_transform_pcd(self.h, _p(xyzi, _f32p), len(xyzi), _p(T, _f64p), _p(out, _f32p)))
        return out

    def voxelize_pcd(self, xyzi, voxel_res=None):
        xyzi = _f32(xyzi).reshape(-1, 4)
        out = np.zeros_like(xyzi)
        n = C.c_int(0)
        self._check(self.lib.loop_voxelize(self.h, _p(xyzi, _f32p), len(xyzi), self.voxel_res if voxel_res is None else voxel_res,
                                           _p(out, _f32p), C.byref(n)))
        return out[: n.value].copy()

    # ---- measurement hooks
    def timing(self, on=True):
        self.lib.loop_timing_enable(self.h, int(on))

    def timing_reset(self):
        self.lib.loop_timing_reset(self.h)

    def timing_get(self, name):
        ms, n = C.c_double(0), C.c_int(0)
        self._check(self.lib.loop_timing_get(self.h, name.encode(), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def launch_count(self):
        return int(self.lib.loop_launch_count(self.h))

    def batch_stats(self):
        a, b, c = C.c_int64(0), C.c_int64(0), C.c_int64(0)
        self._check(self.lib.loop_batch_stats(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(iterations=a.value, src_points=b.value, nn_queries=c.value)



class LoopClosurePool:
    """Several LoopClosure contexts on ONE GPU (own stream and workspace each, the keyframe store replicated), driven from
    host threads: a batch of candidate pairs is cut into contiguous parts that are verified concurrently.  Candidate pairs are
    independent (loop_closure.cpp:101-126), so the results are those of a single context; what the pool buys is overlap - the
    short, latency-bound stretches of one part (VoxelGrid bookkeeping, tree build, the last ICP iterations of a few straggler
    pairs, host round trips) run under the wide correspondence-search kernels of the other.  That matters most when a rank
    holds few pairs (8-GPU sharding of a 256-pair batch)."""

    def __init__(self, n_ctx=2, **kw):
        self.ctx = [LoopClosure(**kw) for _ in range(max(1, int(n_ctx)))]

    def close(self):
        for c in self.ctx:
            c.close()

    def add_keyframe(self, xyzi, pose_corrected, stamp=0.0):
        idx = -1
        for c in self.ctx:
            idx = c.add_keyframe(xyzi, pose_corrected, stamp)
        return idx

    def clear(self):
        for c in self.ctx:
            c.clear()

    def split(self, n):
        """[lo, hi) of each context's contiguous part of an n-pair list."""
        return [shard_pairs(n, k, len(self.ctx)) for k in range(len(self.ctx))]

    def make_arrays(self, pairs):
        """ctypes pair / result arrays, one set per context (caller-owned, reusable across calls)."""
        out = []
        for lo, hi in self.split(len(pairs)):
            n = hi - lo
            arr = (LoopPair * max(n, 1))()
            for i, (a, b) in enumerate(pairs[lo:hi]):
                arr[i].src_idx, arr[i].dst_idx = int(a), int(b)
            out.append((arr, n, (LoopResult * max(n, 1))()))
        return out

    def verify_arrays(self, arrays):
        """One loop_verify_batch per context, concurrently (the C call releases the GIL and ends with its stream synchronised)."""
        import threading

        errs = []

        def run(c, a):
            try:
                if a[1]:
                    c.verify_batch_raw(a[0], a[1], a[2])
            except Exception as e:  # surfaced below: a failed part must fail the batch
                errs.append(e)

        th = [threading.Thread(target=run, args=(c, a)) for c, a in zip(self.ctx[1:], arrays[1:])]
        for t in th:
            t.start()
        run(self.ctx[0], arrays[0])
        for t in th:
            t.join()
        if errs:
            raise errs[0]

    def verify_batch(self, pairs):
        arrays = self.make_arrays(pairs)
        self.verify_arrays(arrays)
        return [_result(res[i]) for arr, n, res in arrays for i in range(n)]
