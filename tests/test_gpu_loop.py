"""GPU parity tests of the loop-closure verification path, called through the C-ABI (include/loop_b200.h).

Bar: everything on this path is compared BIT-EXACT with the CPU oracle on identical seeded inputs — voxelised submaps
(point sets and order), ICP correspondences, final transforms, iteration counts, stop reasons and fitness scores —
because the reductions are exact fixed-point sums and the f32 operation order is shared (oracle/orc_icp.cpp).
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def loop_mod(pkg):
    import importlib

    return importlib.import_module("fast-lio-sam_gps_b200.loop")


@pytest.fixture(scope="module")
def drive(synth):
    scene = synth.make_scene(half_extent=60.0)
    return synth.make_keyframes(scene, n_pass=14, n_az=360, max_range=50.0)


@pytest.fixture(scope="module")
def closers(loop_mod, orc, drive):
    clouds, est, _, stamps = drive
    g = loop_mod.LoopClosure(num_submap_keyframes=3)
    o = orc.OracleLoop(max_corr_dist=g.icp_max_corr_dist, nthreads=0)
    for c, e, s in zip(clouds, est, stamps):
        assert g.add_keyframe(c, e, s) == o.add_keyframe(c, e, s)
    yield g, o
    g.close()


def same_result(r, ref, exact=True):
    assert r["iterations"] == ref["iterations"] and r["converged"] == ref["converged"]
    if "state" in ref:
        assert r["state"] == ref["state"]
    assert np.array_equal(r["T"], ref["T"])
    assert r["score"] == ref["score"]


def test_transform_pcd_bit_exact(closers, orc, synth, drive):
    g, _ = closers
    T = synth.pose44(synth.rot_from_rpy(0.1, -0.2, 0.7), [10.0, -20.0, 3.0])
    assert np.array_equal(g.transform_pcd(drive[0][3], T), orc.transform_pcd(drive[0][3], T))
    assert len(g.transform_pcd(np.zeros((0, 4), np.float32), T)) == 0


def test_keyframe_add_world_bit_exact(loop_mod, orc, synth, drive):
    """loop_keyframe_add_world = PosePcd's constructor on the device (pose_pcd.hpp:37-39): the stored cloud equals
    transformPcd(cloud_world, pose.inverse()) of the oracle bit for bit, and a submap built from such keyframes equals the one
    built from keyframes ingested in the LiDAR frame by the oracle's constructor."""
    clouds, est, _, stamps = drive
    g = loop_mod.LoopClosure(num_submap_keyframes=2)
    o = orc.OracleLoop(max_corr_dist=g.icp_max_corr_dist, nthreads=0)
    for k in range(8):
        world = orc.transform_pcd(clouds[k], est[k])  # what FAST-LIO publishes on /cloud_registered
        assert g.add_keyframe_world(world, est[k], stamps[k]) == k
        o.add_keyframe(orc.keyframe_from_world(world, est[k]), est[k], stamps[k])
    assert np.abs(orc.pose_inverse(est[3]) - np.linalg.inv(np.asarray(est[3], np.float64))).max() < 1e-12
    gs, gd = g.set_src_and_dst_cloud(5, 2, 2, 0.3)
    os_, od = o.set_src_dst(5, 2, 2, 0.3)
    assert np.array_equal(gs, os_) and np.array_equal(gd, od)
    with pytest.raises(Exception):
        g.add_keyframe_world(clouds[0], np.zeros((4, 4)), 0.0)  # singular pose
    g.close()


def test_fetch_closest_batch_matches(closers, drive):
    """loop_fetch_closest_batch (one device launch) = loop_fetch_closest per query = the reference's scan
    (loop_closure.cpp:18-40), before and after a bulk pose update (loop_keyframe_set_poses)."""
    g, o = closers
    K = len(drive[0])
    q = np.arange(K, dtype=np.int32)
    stamps = np.asarray(drive[3], np.float64)
    keep = g.loop_detection_timediff_threshold
    g.loop_detection_timediff_threshold = 0.25 * float(stamps[-1] - stamps[0])  # so that the return pass finds the outbound one
    try:
        _fetch_closest_checks(g, o, drive, K, q)
    finally:
        g.loop_detection_timediff_threshold = keep


def _fetch_closest_checks(g, o, drive, K, q):
    one = np.array([g.fetch_closest_keyframe_idx(int(i)) for i in q], np.int32)
    assert np.array_equal(g.fetch_closest_keyframe_idx_batch(q), one)
    assert np.array_equal(one, np.array([o.fetch_closest(int(i), g.loop_detection_radius, g.loop_detection_timediff_threshold) for i in q]))
    assert (one >= 0).any() and (one < 0).any()
    # move a few poses (as a pose-graph update does), only those are re-sent
    est = [np.asarray(e, np.float64).reshape(4, 4).copy() for e in drive[1]]
    idx = np.array([1, 4, K - 3], np.int32)
    moved = []
    for i in idx:
        T = est[i].copy()
        T[:3, 3] += [0.7, -0.4, 0.1]
        moved.append(T)
    g.set_poses(idx, np.array(moved))
    for i, T in zip(idx, moved):
        o.set_pose(int(i), T)
    again = g.fetch_closest_keyframe_idx_batch(q)
    assert np.array_equal(again, np.array([g.fetch_closest_keyframe_idx(int(i)) for i in q], np.int32))
    assert np.array_equal(again, np.array([o.fetch_closest(int(i), g.loop_detection_radius, g.loop_detection_timediff_threshold) for i in q]))
    g.set_poses(idx, np.array([est[i] for i in idx]))  # restore for the tests that follow
    for i in idx:
        o.set_pose(int(i), est[i])
    assert len(g.fetch_closest_keyframe_idx_batch(np.zeros(0, np.int32))) == 0


def test_pose_pcd_ctor_stores_lidar_frame(closers, orc, synth, drive):
    """pose_pcd.hpp:38-39: the published world-frame cloud is re-expressed in the LiDAR frame with pose^-1; and the loop
    factor built from a verified pair (fast_lio_sam.cpp:699-717) closes the chain pose_from * measured = pose_to."""
    import importlib

    tj = importlib.import_module("fast-lio-sam_gps_b200.trajectory")
    g, _ = closers
    clouds, est = drive[0], drive[1]
    world = g.transform_pcd(clouds[2], est[2])
    kf = tj.PosePcd.from_world_cloud(g, world, est[2], 12.5, 2)
    inv = np.linalg.inv(np.asarray(est[2], np.float64))
    inv[3] = [0, 0, 0, 1]
    assert np.allclose(kf.pcd, orc.transform_pcd(world, np.ascontiguousarray(inv)), atol=1e-5)
    assert np.allclose(kf.pcd[:, :3], clouds[2][:, :3], atol=2e-4) and np.array_equal(kf.pcd[:, 3], clouds[2][:, 3])
    assert kf.idx == 2 and kf.timestamp == 12.5 and np.array_equal(kf.pose_corrected_eig, kf.pose_eig)
    K = len(clouds)
    built = 0
    for q, closest in ((K - 2, 1), (K - 3, 2), (K - 4, 4)):
        r = g.perform_loop_closure(q, closest)
        f = tj.loop_factor(r, est[q], est[closest], q, closest)
        if r["valid"]:
            pose_from = np.asarray(r["T"], np.float64).reshape(4, 4) @ np.asarray(est[q], np.float64)
            # T_icp is a Matrix4f cast to double: its rotation block is orthonormal to f32 rounding only
            assert np.allclose(pose_from @ f["measured"], est[closest], atol=1e-4) and np.all(f["variances"] == r["score"])
            assert (f["key_from"], f["key_to"]) == (q, closest)
            built += 1
        else:
            assert f is None
    assert built >= 1


def test_voxelize_bit_exact(closers, orc, drive):
    g, _ = closers
    rng = np.random.default_rng(1)
    for cloud in (drive[0][0], np.concatenate([rng.uniform(-4, 4, (30000, 3)), rng.uniform(0, 255, (30000, 1))], 1).astype(np.float32)):
        for leaf in (0.3, 1.0, 0.05):
            assert np.array_equal(g.voxelize_pcd(cloud, leaf), orc.voxelize(cloud, leaf))


def test_voxelize_edge_cases(closers, orc):
    g, _ = closers
    assert len(g.voxelize_pcd(np.zeros((0, 4), np.float32), 0.3)) == 0
    one = np.array([[1, 2, 3, 4]], np.float32)
    assert np.array_equal(g.voxelize_pcd(one, 0.3), one)
    assert np.array_equal(g.voxelize_pcd(np.repeat(one, 7, 0), 0.3), one)
    far = np.array([[0, 0, 0, 1], [1e6, 1e6, 1e6, 2], [5, 5, 5, 3]], np.float32)  # PCL's index-overflow guard: unfiltered
    assert np.array_equal(g.voxelize_pcd(far, 0.01), orc.voxelize(far, 0.01))
    assert np.array_equal(g.voxelize_pcd(far, 0.01), far)
    nan = np.array([[0, 0, 0, 1], [np.nan, 0, 0, 2], [0.1, 0.1, 0.1, 3], [np.inf, 1, 1, 4]], np.float32)  # non-finite points are dropped
    assert np.array_equal(g.voxelize_pcd(nan, 0.3), orc.voxelize(nan, 0.3))


def test_fetch_closest_matches(closers, drive):
    g, o = closers
    n = len(drive[0])
    for q in range(n):
        for radius, tdiff in ((35.0, 30.0), (35.0, 5.0), (3.0, 2.0)):
            g.loop_detection_radius, g.loop_detection_timediff_threshold = radius, tdiff
            assert g.fetch_closest_keyframe_idx(q) == o.fetch_closest(q, radius, tdiff)
    g.loop_detection_radius, g.loop_detection_timediff_threshold = 35.0, 30.0


def test_set_src_dst_bit_exact(closers, drive):
    g, o = closers
    n = len(drive[0])
    for (a, b, r) in ((n - 1, 2, 3), (n - 3, 5, 3), (20, 0, 1), (n - 1, n - 1, 2), (15, 3, 0)):
        gs, gd = g.set_src_and_dst_cloud(a, b, r, 0.3)
        os_, od = o.set_src_dst(a, b, r, 0.3)
        assert np.array_equal(gs, os_) and np.array_equal(gd, od)


def test_icp_alignment_bit_exact(closers, drive):
    g, o = closers
    n = len(drive[0])
    src, dst = o.set_src_dst(n - 4, 3, 3, 0.3)
    r = g.icp_alignment(src, dst, want_aligned=True)
    ref = o.icp_align(src, dst, want_aligned=True)
    same_result(r, ref)
    assert np.array_equal(r["corr"], ref["corr"])  # last correspondence search, index by index
    assert np.array_equal(r["aligned"], ref["aligned"])
    assert r["converged"] and 1 <= r["iterations"] <= 50 and r["valid"] == (ref["score"] < 1.5)


def test_icp_edge_cases(loop_mod, orc, closers, drive):
    g, o = closers
    src, dst = o.set_src_dst(len(drive[0]) - 4, 3, 2, 0.3)
    # identical clouds, one iteration allowed: stop reason "iterations", score exactly 0
    g1 = loop_mod.LoopClosure(max_iterations=1)
    o1 = orc.OracleLoop(max_corr_dist=52.5, max_iter=1)
    same_result(g1.icp_alignment(dst, dst), o1.icp_align(dst, dst))
    assert g1.icp_alignment(dst, dst)["score"] == 0.0
    g1.close()
    # fewer than 3 correspondences: not converged, state 5, score still reported (loop_closure.cpp:83), never valid
    g2 = loop_mod.LoopClosure(icp_max_corr_dist=0.01)
    o2 = orc.OracleLoop(max_corr_dist=0.01)
    shifted = dst + np.array([5, 5, 5, 0], np.float32)
    r, ref = g2.icp_alignment(shifted, dst), o2.icp_align(shifted, dst)
    same_result(r, ref)
    assert not r["converged"] and not r["valid"] and r["state"] == 5 and np.isfinite(r["score"])
    # empty source / empty target / tiny clouds
    e = np.zeros((0, 4), np.float32)
    for s_, d_ in ((e, dst), (src, e), (src[:2], dst), (src, dst[:5]), (src[:9], dst[:9])):
        same_result(g2.icp_alignment(s_, d_), o2.icp_align(s_, d_))
        same_result(g.icp_alignment(s_, d_), o.icp_align(s_, d_))
    g2.close()


def test_verify_batch_bit_exact(closers, drive):
    g, o = closers
    n = len(drive[0])
    pairs = [(n - 1 - i, i + (i % 3)) for i in range(1, 13)] + [(n - 2, 1), (n - 2, 1), (5, 5), (n - 1, 0)]
    res = g.verify_batch(pairs)
    assert len(res) == len(pairs)
    valid = 0
    for (a, b), r in zip(pairs, res):
        ref = o.verify(a, b, 3, 0.3, 1.5)
        assert (r["n_src"], r["n_dst"]) == (ref["n_src"], ref["n_dst"])
        same_result(r, ref)
        assert r["valid"] == ref["valid"]
        valid += r["valid"]
    assert valid >= 8
    st = g.batch_stats()
    assert st["iterations"] == sum(r["iterations"] for r in res) and st["src_points"] == sum(r["n_src"] for r in res)
    # debug getters of the last batch (src_cloud_, dst_cloud_, aligned_)
    k = 3
    os_, od = o.set_src_dst(pairs[k][0], pairs[k][1], 3, 0.3)
    assert np.array_equal(g.get_cloud(k, "src"), os_) and np.array_equal(g.get_cloud(k, "dst"), od)
    ref = o.icp_align(os_, od, want_aligned=True)
    assert np.array_equal(g.get_cloud(k, "aligned"), ref["aligned"])
    # a single candidate through the per-call mirror, and "no candidate"
    one = g.perform_loop_closure(pairs[0][0], pairs[0][1])
    assert np.array_equal(one["T"], res[0]["T"]) and one["score"] == res[0]["score"]
    assert g.verify_batch([]) == []


def test_batch_is_order_and_chunk_independent(closers, drive):
    g, _ = closers
    n = len(drive[0])
    pairs = [(n - 1 - i, i) for i in range(1, 9)]
    a = g.verify_batch(pairs)
    b = g.verify_batch(pairs[::-1])[::-1]
    for x, y in zip(a, b):
        assert np.array_equal(x["T"], y["T"]) and x["score"] == y["score"] and x["iterations"] == y["iterations"]


def test_pose_update_changes_submap(closers, synth, drive):
    g, o = closers
    clouds, est, _, _ = drive
    T = est[4].copy()
    T[:3, 3] += [0.2, 0.0, 0.0]
    g.set_pose(4, T)
    gs, _ = g.set_src_and_dst_cloud(4, 4, 0, 0.3)
    assert np.array_equal(gs, orc_vox(o, clouds[4], T))
    g.set_pose(4, est[4])


def orc_vox(o, cloud, T):
    from oracle import orc

    return orc.voxelize(orc.transform_pcd(cloud, T), 0.3)


def test_golden_loop_gpu(loop_mod):
    gld = np.load(os.path.join(GOLD, "loop_small.npz"))
    g = loop_mod.LoopClosure(num_submap_keyframes=1)
    for i in range(6):
        Tp = gld["poses"][i].copy()
        if i >= 3:
            Tp[:3, 3] += [0.3, -0.2, 0.05]
        g.add_keyframe(gld[f"kf{i}"], Tp, stamp=float(i))
    src, dst = g.set_src_and_dst_cloud(4, 1, 1, 0.3)
    assert np.array_equal(src, gld["src"]) and np.array_equal(dst, gld["dst"])
    r = g.icp_alignment(src, dst, want_aligned=True)
    assert r["iterations"] == int(gld["iterations"]) and r["state"] == int(gld["state"]) and r["converged"] == bool(gld["converged"])
    assert np.array_equal(r["T"], gld["T"]) and r["score"] == float(gld["score"])
    assert np.array_equal(r["aligned"], gld["aligned"])
    g.close()


def test_errors(loop_mod, pkg):
    g = loop_mod.LoopClosure()
    with pytest.raises(pkg.LioError):
        g.verify_batch([(0, 1)])  # empty keyframe store
    with pytest.raises(pkg.LioError):
        g.fetch_closest_keyframe_idx(0)
    g.add_keyframe(np.zeros((0, 4), np.float32), np.eye(4))  # empty keyframes are legal
    g.add_keyframe(np.zeros((0, 4), np.float32), np.eye(4))
    r = g.verify_batch([(1, 0)])[0]
    assert not r["converged"] and not r["valid"] and r["n_src"] == 0 and r["n_dst"] == 0 and r["state"] == 5
    with pytest.raises(pkg.LioError):
        g.get_cloud(7, "src")
    g.close()
    with pytest.raises(pkg.LioError):
        loop_mod.LoopClosure(max_iterations=0)


def test_pool_equals_single_context(loop_mod, closers, drive):
    """LoopClosurePool: a batch cut into contiguous parts verified concurrently by several contexts gives the results of one
    context, pair for pair (candidate pairs are independent, loop_closure.cpp:101-126)."""
    g, _ = closers
    clouds, est, _, stamps = drive
    K = len(clouds)
    pairs = [(K - 2 - i, 1 + (i % 4)) for i in range(7)]
    ref = g.verify_batch(pairs)
    for n_ctx in (2, 3):
        pool = loop_mod.LoopClosurePool(n_ctx, num_submap_keyframes=g.num_submap_keyframes)
        for c, e, s in zip(clouds, est, stamps):
            pool.add_keyframe(c, e, s)
        assert [hi - lo for lo, hi in pool.split(len(pairs))] == [len(range(*loop_mod.shard_pairs(len(pairs), k, n_ctx))) for k in range(n_ctx)]
        got = pool.verify_batch(pairs)
        assert len(got) == len(ref)
        for a, b in zip(got, ref):
            same_result(a, b)
        pool.close()


@pytest.mark.parametrize("knob", [("LOOP_TREE", "balanced"), ("LOOP_GRID", "1"), ("LOOP_SAMPLE_FIRST", "0"), ("LOOP_TREE_ORDER", "morton")])
def test_search_knobs_change_nothing(loop_mod, closers, drive, monkeypatch, knob):
    """The A/B switches of the correspondence search (index-balanced tree, grid probe for seeded searches, cold first
    iteration, Morton order) only change how the exact nearest neighbour is found: every result is bit-identical."""
    g, _ = closers
    clouds, est, _, stamps = drive
    K = len(clouds)
    pairs = [(K - 2 - i, 1 + (i % 3)) for i in range(5)]
    ref = g.verify_batch(pairs)
    monkeypatch.setenv(*knob)
    h = loop_mod.LoopClosure(num_submap_keyframes=g.num_submap_keyframes)
    for c, e, s in zip(clouds, est, stamps):
        h.add_keyframe(c, e, s)
    got = h.verify_batch(pairs)
    for a, b in zip(got, ref):
        same_result(a, b)
    h.close()
