"""GPU parity tests of the IEKF measurement path, called through the C-ABI (include/lio_b200.h).

Bars (BASELINE.json north_star): kNN index sets and plane-validity masks bit-exact; H^T H / H^T z within
1e-5 relative; per-iteration state updates within 1e-6 m / 1e-6 rad; trajectory divergence < 1 mm.
The checker is the CPU oracle on identical seeded inputs.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL = 1e-5  # north_star tolerance on H^T H / H^T z
STATE_TOL = 1e-6  # m / rad per iteration


def rel_err(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.fixture(scope="module")
def pair(pkg, orc, small_world):
    g = pkg.LioContext(0.5, max_points=1 << 19, max_scan=1 << 16)
    o = orc.OracleLio(0.5, nthreads=0)
    g.map_build(small_world["map"])
    o.map_build(small_world["map"])
    yield g, o
    g.close()
    o.close()


def test_library_is_the_cuda_one(pkg):
    lib = pkg.load_library()
    assert b"sm_100a" in lib.lio_version()
    assert lib.lio_device_count() >= 1


def test_map_build_roundtrip(pair, small_world):
    g, o = pair
    assert g.map_size() == o.map_size() == len(small_world["map"])
    xyz, ids = g.map_export()
    assert np.array_equal(ids, np.arange(len(ids), dtype=np.uint32))
    assert np.array_equal(xyz, small_world["map"])


def test_knn_bit_exact(pair, small_world):
    g, o = pair
    rng = np.random.default_rng(7)
    m = small_world["map"]
    q = m[rng.integers(0, len(m), 20000)] + rng.normal(0, 0.4, (20000, 3)).astype(np.float32)
    q = np.concatenate([q, rng.uniform(-90, 90, (2000, 3)).astype(np.float32), m[:2000]])  # far, and exact hits (d2 = 0)
    for k in (5, 1, 3):
        gi, gd, gp, gc = g.map_knn(q, k, 5.0)
        oi, od, op, oc = o.map_knn(q, k, 5.0)
        assert np.array_equal(gc, oc)
        assert np.array_equal(gi, oi)
        assert np.array_equal(gd, od)
        assert np.array_equal(gp, op)
    assert (gc < 3).any() and (gc == 3).any()  # both padded and full rows were exercised


def test_scan_prefetch_equals_plain_upload(pair, small_world):
    """lio_scan_prefetch + lio_scan_upload (copy on the copy stream, double-buffered) feeds the filter the same scan as a
    plain lio_scan_upload; an upload from another pointer ignores a pending prefetch."""
    g, _ = pair
    st = small_world["start"]
    a = np.ascontiguousarray(small_world["scan"], np.float32)
    b = np.ascontiguousarray(a[::-1] + np.float32(0.01))
    ref = {}
    for name, s in (("a", a), ("b", b)):
        g.scan_upload(s)
        ref[name] = g.h_share(st, True)
    for rnd in range(3):  # alternate slots: a, b, a through the prefetch path
        name, s = (("a", a), ("b", b))[rnd % 2]
        g.scan_prefetch_ptr(s.ctypes.data, len(s))
        g.scan_upload_ptr(s.ctypes.data, len(s))
        H, z, n_eff, res = g.h_share(st, True)
        assert n_eff == ref[name][2] and np.array_equal(H, ref[name][0]) and np.array_equal(z, ref[name][1])
    g.scan_prefetch_ptr(b.ctypes.data, len(b))  # pending prefetch of b, but a is what gets uploaded
    g.scan_upload_ptr(a.ctypes.data, len(a))
    H, z, n_eff, res = g.h_share(st, True)
    assert n_eff == ref["a"][2] and np.array_equal(H, ref["a"][0])
    g.scan_upload_ptr(b.ctypes.data, len(b))  # the stale prefetch of b is still valid for b's pointer and is consumed here
    H, z, n_eff, res = g.h_share(st, True)
    assert n_eff == ref["b"][2] and np.array_equal(H, ref["b"][0])


@pytest.mark.parametrize("left", ["group", "warp"])
def test_knn_ring_expansion_is_exact(pkg, orc, left, monkeypatch):
    """Sparse map + large radius: the answer lies several hash cells away, so rings 2.. must be searched.  Both forms of
    the leftover kernel: 8-lane groups (generic batches) and warp per query with the occupancy words (the scan path's)."""
    monkeypatch.setenv("LIO_BATCH_LEFT", left)
    rng = np.random.default_rng(8)
    m = rng.uniform(-20, 20, (3000, 3)).astype(np.float32)
    q = rng.uniform(-22, 22, (4000, 3)).astype(np.float32)
    g = pkg.LioContext(0.5, max_points=1 << 14, max_scan=1024)
    o = orc.OracleLio(0.5)
    g.map_build(m)
    o.map_build(m)
    for r2 in (5.0, 30.0, 100.0):
        gi, gd, _, gc = g.map_knn(q, 5, r2)
        oi, od, _, oc = o.map_knn(q, 5, r2)
        assert np.array_equal(gi, oi) and np.array_equal(gd, od) and np.array_equal(gc, oc)
    g.close()


def test_knn_ties_and_duplicates(pkg, orc):
    pts = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1], [3, 0, 0], [1, 0, 0], [1, 0, 0]], np.float32)
    g = pkg.LioContext(0.5, max_points=1024, max_scan=128)
    o = orc.OracleLio(0.5)
    g.map_build(pts)
    o.map_build(pts)
    q = np.array([[0, 0, 0], [1, 0, 0], [0.5, 0, 0]], np.float32)
    gi, gd, _, _ = g.map_knn(q, 5, 5.0)
    oi, od, _, _ = o.map_knn(q, 5, 5.0)
    assert np.array_equal(gi, oi) and np.array_equal(gd, od)
    assert gi[0].tolist() == [0, 1, 2, 3, 4]
    g.close()


@pytest.mark.parametrize("ext", [False, True])
def test_h_share_parity(pkg, orc, small_world, ext):
    w = small_world
    g = pkg.LioContext(0.5, max_points=1 << 19, max_scan=1 << 16, extrinsic_est_en=ext)
    o = orc.OracleLio(0.5, nthreads=0, extrinsic_est_en=ext)
    g.map_build(w["map"])
    o.map_build(w["map"])
    g.scan_upload(w["scan"])
    o.scan_upload(w["scan"])
    st = w["start"].copy()
    if ext:
        st[21:24] = [0.05, -0.02, 0.1]
    prev_sel = None
    for redo, state in ((True, st), (False, w["truth"] if not ext else st), (False, st), (True, w["truth"])):
        gH, gz, gn, gr = g.h_share(state, redo)
        oH, oz, on, orr = o.h_share(state, redo)
        go, oo = g.scan_get(), o.scan_get()
        if not redo:  # point_selected_surf persists between searches: the set can only shrink
            assert (go["sel"] <= prev_sel).all()
        prev_sel = go["sel"].copy()
        assert np.array_equal(go["world"], oo["world"])
        assert np.array_equal(go["nn_cnt"], oo["nn_cnt"])
        assert np.array_equal(go["nn_ids"], oo["nn_ids"])  # kNN index sets: bit-exact
        assert np.array_equal(go["nn_xyz"], oo["nn_xyz"])
        assert np.array_equal(go["sel"], oo["sel"])  # plane-validity mask: bit-exact
        s = oo["sel"] == 1
        assert np.array_equal(go["plane"][s], oo["plane"][s])
        assert gn == on and on > 1000
        assert rel_err(gH, oH) < REL and rel_err(gz, oz) < REL
        assert rel_err(gH, oH) < 1e-11  # same f64 rows, only the summation order differs
        assert abs(gr - orr) <= 1e-9 * orr
        assert np.array_equal(gH, gH.T)
    g.close()


def test_point_selected_surf_persists_gpu(pkg, orc, planted):
    """A point that fails the weight test in call 1 stays unselected in the calls that reuse the cached neighbours, although
    it would pass at the new state (the oracle-side statement: test_oracle_lio.py::test_point_selected_surf_persists...)."""
    m, scan, truth, off = planted
    g = pkg.LioContext(0.5, max_points=1 << 16, max_scan=1 << 12)
    o = orc.OracleLio(0.5, nthreads=0)
    g.map_build(m)
    o.map_build(m)
    g.scan_upload(scan)
    o.scan_upload(scan)
    near = np.sqrt(np.linalg.norm(scan, axis=1)) < 1.33
    sels = []
    for redo, state in ((True, off), (False, truth), (True, truth), (False, off), (False, truth)):
        gH, gz, gn, _ = g.h_share(state, redo)
        oH, oz, on, _ = o.h_share(state, redo)
        go, oo = g.scan_get(), o.scan_get()
        assert np.array_equal(go["sel"], oo["sel"]) and gn == on == go["sel"].sum()
        assert rel_err(gH, oH) < 1e-11 and rel_err(gz, oz) < 1e-9
        sels.append(go["sel"].copy())
    assert sels[0][near].sum() == 0 and sels[1][near].sum() == 0 and np.array_equal(sels[0], sels[1])
    assert sels[2].all()
    assert sels[3][near].sum() == 0 and np.array_equal(sels[3], sels[4])
    # the rows entry point and the device-resident loop go through the same kernel: same flag
    sel_r, rows, z, n_r = g.h_share_rows(truth, False)
    assert np.array_equal(sel_r, sels[4]) and n_r == sels[4].sum()
    g.close()


def test_h_share_rows_fill_dyn_share(pair, small_world):
    """lio_h_share_rows delivers what an adapter needs to fill the unchanged dyn_share_datastruct (h_x, h)."""
    g, o = pair
    w = small_world
    g.scan_upload(w["scan"])
    o.scan_upload(w["scan"])
    sel, rows, z, n_eff = g.h_share_rows(w["start"], True)
    oH, oz, on, _, hx, hz = o.h_share(w["start"], True, want_rows=True)
    assert n_eff == on == sel.sum()
    assert np.allclose(rows[sel == 1], hx, rtol=0, atol=1e-12)
    assert np.allclose(z[sel == 1], hz, rtol=0, atol=0)
    assert (rows[sel == 0] == 0).all()


def test_iekf_update_parity(pair, synth, small_world):
    g, o = pair
    w = small_world
    g.scan_upload(w["scan"])
    o.scan_upload(w["scan"])
    P0 = synth.default_cov()
    gs, gP, gi = g.iekf_update(w["start"], P0, trace=True)
    os_, oP, oi = o.iekf_update(w["start"], P0, trace=True)
    assert gi["iters"] == oi["iters"]
    assert np.array_equal(gi["n_eff"], oi["n_eff"]) and np.array_equal(gi["knn"], oi["knn"])
    assert np.abs(gi["states"] - oi["states"]).max() < STATE_TOL  # per-iteration state updates
    assert np.abs(gs - os_).max() < STATE_TOL
    assert np.abs(gP - oP).max() < 1e-9
    assert np.linalg.norm(gs[:3] - w["truth"][:3]) < 0.01


def _ctx(pkg, mode, **kw):
    """A context whose lio_iekf_update runs the device-resident loop ('dev') or the host loop over lio_h_share ('host')."""
    old = os.environ.get("LIO_IEKF_MODE")
    os.environ["LIO_IEKF_MODE"] = "host" if mode == "host" else "device"
    try:
        return pkg.LioContext(0.5, **kw)
    finally:
        if old is None:
            del os.environ["LIO_IEKF_MODE"]
        else:
            os.environ["LIO_IEKF_MODE"] = old


@pytest.mark.gpu
@pytest.mark.parametrize("ext", [False, True])
def test_device_loop_equals_host_loop(pkg, synth, small_world, ext):
    """The device-resident IEKF loop (lio_iekf.cuh) and the host loop (esekf_host.cpp) run the same FP64 operations in the
    same order; only libm (sin/cos/atan2) may differ in the last bits."""
    w = small_world
    P0 = synth.default_cov()
    out = {}
    for mode in ("dev", "host"):
        g = _ctx(pkg, mode, max_points=1 << 19, max_scan=1 << 16, extrinsic_est_en=ext)
        g.map_build(w["map"])
        g.scan_upload(w["scan"])
        out[mode] = g.iekf_update(w["start"], P0, trace=True)
        out[mode + "_scan"] = g.scan_get()
        g.close()
    (ds, dP, di), (hs, hP, hi) = out["dev"], out["host"]
    assert di["iters"] == hi["iters"] and np.array_equal(di["n_eff"], hi["n_eff"]) and np.array_equal(di["knn"], hi["knn"])
    assert np.abs(di["states"] - hi["states"]).max() < 1e-11
    assert np.abs(ds - hs).max() < 1e-11 and np.abs(dP - hP).max() < 1e-12
    for k in ("nn_ids", "sel"):
        assert np.array_equal(out["dev_scan"][k], out["host_scan"][k])


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["host", "dev"])
def test_iekf_update_map_parity(pkg, orc, synth, small_world, mode):
    """lio_iekf_update_map = update_iterated_dyn_share_modified + map_incremental in one enqueue: same state, covariance
    and live map as the oracle's two calls; three scans in a row so the second search runs on the grown map."""
    w = small_world
    g = _ctx(pkg, mode, max_points=1 << 19, max_scan=1 << 16)
    o = orc.OracleLio(0.5)
    g.map_build(w["map"])
    o.map_build(w["map"])
    P0 = synth.default_cov()
    rng = np.random.default_rng(5)
    for rnd in range(3):
        scan = (w["scan"] + rng.normal(0, 0.01, w["scan"].shape)).astype(np.float32)
        g.scan_upload(scan)
        o.scan_upload(scan)
        gs, gP, gi = g.iekf_update_map(w["start"], P0, True, trace=True)
        os_, oP, oi = o.iekf_update(w["start"], P0, trace=True)
        o.map_incremental(os_, True)
        assert gi["iters"] == oi["iters"] and np.array_equal(gi["n_eff"], oi["n_eff"]) and np.array_equal(gi["knn"], oi["knn"])
        assert np.abs(gi["states"] - oi["states"]).max() < STATE_TOL and np.abs(gP - oP).max() < 1e-9
        gx, gid = g.map_export()
        ox, oid = o.map_export()
        assert g.map_size() == o.map_size()
        assert np.array_equal(gid, oid) and np.array_equal(gx, ox)
    g.close()


@pytest.mark.gpu
def test_graphs_serve_changing_scan_sizes(pkg, orc, synth, small_world):
    """The captured graphs (scan ordering, map_incremental) are cached per padded size and read the true size from device
    memory: scans of different lengths through ONE context must match the oracle and a graph-free context bit for bit."""
    w = small_world
    P0 = synth.default_cov()
    old = {k: os.environ.get(k) for k in ("LIO_SORT_GRAPH", "LIO_MAP_GRAPH")}
    try:
        g = pkg.LioContext(0.5, max_points=1 << 19, max_scan=1 << 16)
        os.environ["LIO_SORT_GRAPH"] = os.environ["LIO_MAP_GRAPH"] = "0"
        p = pkg.LioContext(0.5, max_points=1 << 19, max_scan=1 << 16)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    o = orc.OracleLio(0.5)
    for c in (g, p, o):
        c.map_build(w["map"])
    n = len(w["scan"])
    rng = np.random.default_rng(11)
    for m in (n, n - 1000, 8192, 8193, 777, n):  # same bucket, other buckets, bucket edges, tiny, back again
        scan = np.ascontiguousarray(w["scan"][rng.permutation(n)[:m]])
        outs = []
        for c in (g, p):
            c.scan_upload(scan)
            outs.append(c.iekf_update_map(w["start"], P0, True, trace=True))
        o.scan_upload(scan)
        os_, oP, oi = o.iekf_update(w["start"], P0, trace=True)
        o.map_incremental(os_, True)
        (gs, gP, gi), (ps, pP, pi) = outs
        assert np.array_equal(gs, ps) and np.array_equal(gP, pP) and np.array_equal(gi["n_eff"], pi["n_eff"])
        assert np.array_equal(gi["n_eff"], oi["n_eff"]) and np.abs(gs - os_).max() < STATE_TOL
        gx, gid = g.map_export()
        px, pid = p.map_export()
        ox, oid = o.map_export()
        assert np.array_equal(gid, oid) and np.array_equal(gx, ox) and np.array_equal(pid, oid) and np.array_equal(px, ox)
    g.close()
    p.close()


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["host", "dev"])
def test_iekf_invalid_rounds_and_errors(pkg, orc, synth, small_world, mode):
    w = small_world
    P0 = synth.default_cov()
    # no effective points at all (scan far away from the map): every round invalid, state and covariance unchanged
    g = _ctx(pkg, mode, max_points=1 << 19, max_scan=1 << 16)
    o = orc.OracleLio(0.5)
    g.map_build(w["map"])
    o.map_build(w["map"])
    far = (w["scan"] + np.float32(5000.0)).astype(np.float32)
    g.scan_upload(far)
    o.scan_upload(far)
    gs, gP, gi = g.iekf_update(w["start"], P0, trace=True)
    os_, oP, oi = o.iekf_update(w["start"], P0, trace=True)
    assert gi["iters"] == oi["iters"] == 4 and not gi["n_eff"].any()
    assert np.array_equal(gs, w["start"]) and np.array_equal(gP, P0) and np.array_equal(os_, gs)
    # a singular covariance is reported, not propagated
    g.scan_upload(w["scan"])
    with pytest.raises(pkg.LioError):
        g.iekf_update(w["start"], np.zeros((23, 23)))
    # and the context stays usable
    gs2, _, _ = g.iekf_update(w["start"], P0)
    o.scan_upload(w["scan"])
    os2, _, _ = o.iekf_update(w["start"], P0)
    assert np.abs(gs2 - os2).max() < STATE_TOL
    # without a scan: call-order error, as on the host path
    g2 = _ctx(pkg, mode, max_points=1 << 12, max_scan=64)
    with pytest.raises(pkg.LioError):
        g2.iekf_update(w["start"], P0)
    g2.close()
    g.close()


def test_map_add_downsample_parity(pkg, orc):
    rng = np.random.default_rng(9)
    g = pkg.LioContext(0.5, max_points=1 << 16, max_scan=1024)
    o = orc.OracleLio(0.5)
    base = rng.uniform(-6, 6, (5000, 3)).astype(np.float32)
    g.map_build(base)
    o.map_build(base)
    for rnd in range(4):
        # dense batches: many points per voxel, voxels that already hold several points, exact duplicates
        batch = rng.uniform(-6.5, 6.5, (4000, 3)).astype(np.float32)
        batch[100:110] = batch[0:10]
        down = rnd != 2
        gnet = g.map_add(batch, down)
        onet, _ = o.map_add(batch, down)
        assert gnet == onet
        gx, gi = g.map_export()
        ox, oi = o.map_export()
        assert g.map_size() == o.map_size() == len(oi)
        assert np.array_equal(gi, oi) and np.array_equal(gx, ox)
    q = rng.uniform(-6, 6, (3000, 3)).astype(np.float32)
    a, b = g.map_knn(q, 5, 5.0), o.map_knn(q, 5, 5.0)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    g.close()


def test_map_delete_boxes_parity(pkg, orc):
    rng = np.random.default_rng(10)
    g = pkg.LioContext(0.5, max_points=1 << 16, max_scan=1024)
    o = orc.OracleLio(0.5)
    base = rng.uniform(-8, 8, (20000, 3)).astype(np.float32)
    g.map_build(base)
    o.map_build(base)
    boxes = np.array([[-2, -2, -2, 1, 3, 2], [4, 4, 4, 9, 9, 9], [-9, -9, -9, -8.5, 9, 9]], np.float32)
    assert g.map_delete_boxes(boxes) == o.map_delete_boxes(boxes) > 0
    assert g.map_delete_boxes(boxes) == 0
    gx, gi = g.map_export()
    ox, oi = o.map_export()
    assert np.array_equal(gi, oi) and np.array_equal(gx, ox)
    # freed slots are reused by later inserts and searches stay exact
    more = rng.uniform(-3, 3, (3000, 3)).astype(np.float32)
    assert g.map_add(more, True) == o.map_add(more, True)[0]
    q = rng.uniform(-8, 8, (3000, 3)).astype(np.float32)
    a, b = g.map_knn(q, 5, 5.0), o.map_knn(q, 5, 5.0)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    g.close()


def test_map_compact_keeps_every_answer(pkg, orc):
    """lio_map_compact: cells / buckets emptied by box deletes are handed back, every search and the flattened map stay
    bit-identical (ids kept); with renumbering the ids become ranks and the neighbours' coordinates stay the same."""
    rng = np.random.default_rng(12)
    g = pkg.LioContext(0.5, max_points=1 << 16, max_scan=1024)
    o = orc.OracleLio(0.5)
    base = rng.uniform(-8, 8, (20000, 3)).astype(np.float32)
    g.map_build(base)
    o.map_build(base)
    boxes = np.array([[-8, -8, -8, 0, 8, 8]], np.float32)  # half of the volume goes
    assert g.map_delete_boxes(boxes) == o.map_delete_boxes(boxes) > 0
    q = rng.uniform(-8, 8, (3000, 3)).astype(np.float32)
    before, st0 = g.map_knn(q, 5, 5.0), g.map_stats()
    g.map_compact()
    st1 = g.map_stats()
    assert st1["live"] == st0["live"] == o.map_size() and st1["next_id"] == st0["next_id"]
    assert st1["n_cells"] < st0["n_cells"] and st1["pool_top"] < st0["pool_top"] and st1["dead_since_compact"] == 0
    after, ref = g.map_knn(q, 5, 5.0), o.map_knn(q, 5, 5.0)
    for a, b, c in zip(before, after, ref):
        assert np.array_equal(a, b) and np.array_equal(b, c)
    gx, gi = g.map_export()
    ox, oi = o.map_export()
    assert np.array_equal(gi, oi) and np.array_equal(gx, ox)
    # inserts after the compaction continue the id sequence exactly as the oracle's
    more = rng.uniform(-3, 3, (3000, 3)).astype(np.float32)
    assert g.map_add(more, True) == o.map_add(more, True)[0]
    gx, gi = g.map_export()
    ox, oi = o.map_export()
    assert np.array_equal(gi, oi) and np.array_equal(gx, ox)
    # renumbering: ids -> ranks, coordinates and order unchanged, the insertion counter restarts at the live count
    g.map_compact(renumber=True)
    gx2, gi2 = g.map_export()
    assert np.array_equal(gx2, gx) and np.array_equal(gi2, np.arange(len(gi), dtype=np.uint32)) and g.map_stats()["next_id"] == len(gi)
    a, b = g.map_knn(q, 5, 5.0), o.map_knn(q, 5, 5.0)
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])  # distances and points
    rank = {int(v): k for k, v in enumerate(oi)}
    assert np.array_equal(a[0], np.vectorize(lambda v: rank.get(int(v), 0xFFFFFFFF), otypes=[np.uint32])(b[0]))
    g.close()


def test_sliding_map_reclaims_space(pkg, orc):
    """A local map that slides (insert ahead, box-delete behind) for far more points than the pools hold: the automatic
    compaction after a delete keeps it running, and the live set stays the oracle's."""
    rng = np.random.default_rng(13)
    g = pkg.LioContext(0.5, max_points=24000, max_cells=8000, max_scan=1024)
    o = orc.OracleLio(0.5)
    total = 0
    for step in range(40):
        x0 = 4.0 * step
        slab = (rng.uniform(0, 1, (6000, 3)) * [4.0, 12.0, 6.0] + [x0, -6.0, -3.0]).astype(np.float32)
        assert g.map_add(slab, True) == o.map_add(slab, True)[0]
        total += len(slab)
        if step >= 2:
            box = np.array([[-1e4, -1e4, -1e4, x0 - 8.0, 1e4, 1e4]], np.float32)
            assert g.map_delete_boxes(box) == o.map_delete_boxes(box)
    st = g.map_stats()
    assert st["error"] == 0 and total > 5 * st["pool_cap"] and st["pool_top"] <= st["pool_cap"] and st["n_cells"] <= st["cell_limit"]
    gx, gi = g.map_export()
    ox, oi = o.map_export()
    assert np.array_equal(gi, oi) and np.array_equal(gx, ox)
    q = (rng.uniform(0, 1, (2000, 3)) * [12.0, 12.0, 6.0] + [4.0 * 37, -6.0, -3.0]).astype(np.float32)
    a, b = g.map_knn(q, 5, 5.0), o.map_knn(q, 5, 5.0)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    g.close()


def test_map_capacity_error_is_loud_and_harmless(pkg, synth, monkeypatch):
    """Pool / table exhaustion: the add reports LIO_E_CAPACITY, no cell ever aliases another cell's bucket, searches keep
    running on what was stored, and the hot path (lio_iekf_update_map) refuses to go on silently."""
    monkeypatch.setenv("LIO_MAP_COMPACT", "0")
    rng = np.random.default_rng(14)
    g = pkg.LioContext(0.5, max_points=2048, max_cells=512, max_scan=4096)
    pts = rng.uniform(-40, 40, (30000, 3)).astype(np.float32)  # ~30 k distinct cells against a limit of a few hundred
    with pytest.raises(pkg.LioError, match="capacity"):
        g.map_add(pts, False)
    st = g.map_stats()
    assert st["error"] != 0 and st["n_cells"] <= st["cell_limit"] and st["live"] <= len(pts)
    # every stored point is still found as its own nearest neighbour (nothing was overwritten through an aliased bucket)
    ids, d2, nn, cnt = g.map_knn(pts[:4000], 1, 1e-12)
    found = cnt == 1
    assert found.sum() > 0 and np.array_equal(nn.reshape(-1, 3)[found], pts[:4000][found])
    scan = rng.uniform(-5, 5, (2000, 3)).astype(np.float32)
    g.scan_upload(scan)
    st0 = synth.make_state(np.eye(3), np.zeros(3))
    with pytest.raises(pkg.LioError, match="capacity"):
        g.iekf_update_map(st0, synth.default_cov(), True)
    g.close()


def test_map_incremental_far_neighbour_parity(pkg, orc, synth):
    """k_far_nn: the scan points whose search finds nothing within max_sqdist are classified with their unbounded nearest map
    point, through the occupancy shells (3-20 m away) and through the full-pool scan (> 32 m): list sizes and the live map
    bit-identical to the oracle, in both launch forms; no search for the scan -> every point is a PointToAdd."""
    from test_oracle_lio import _far_world

    m, scan = _far_world()
    st = synth.make_state(np.eye(3), np.zeros(3))
    for graph in ("1", "0"):
        os.environ["LIO_MAP_GRAPH"] = graph
        g = pkg.LioContext(0.5, max_points=1 << 16, max_scan=4096)
        o = orc.OracleLio(0.5)
        g.map_build(m)
        o.map_build(m)
        g.scan_upload(scan)
        o.scan_upload(scan)
        g.h_share(st, True)
        o.h_share(st, True)
        assert np.array_equal(g.scan_get()["nn_cnt"], o.scan_get()["nn_cnt"])
        got, ref = g.map_incremental(st, True), o.map_incremental(st, True)
        assert got == ref and got[1] > 0
        gx, gi = g.map_export()
        ox, oi = o.map_export()
        assert np.array_equal(gi, oi) and np.array_equal(gx, ox)
        # a scan that was never searched: Nearest_Points is empty for every point
        g.scan_upload(scan + np.float32(0.11))
        o.scan_upload(scan + np.float32(0.11))
        got, ref = g.map_incremental(st, True), o.map_incremental(st, True)
        assert got == ref and got[1] == 0
        gx, gi = g.map_export()
        ox, oi = o.map_export()
        assert np.array_equal(gi, oi) and np.array_equal(gx, ox)
        g.close()
    os.environ.pop("LIO_MAP_GRAPH", None)


def test_golden_map_small_gpu(pkg):
    """The committed map fixture (three scans through map_incremental incl. the far-neighbour branch, one box delete)
    replayed through the C-ABI: list sizes, deleted count and the final live map (ids and coordinates) bit-identical."""
    from test_oracle_lio import _replay_map_small

    g = pkg.LioContext(0.5, max_points=1 << 16, max_scan=4096)
    _replay_map_small(g)
    g.close()


def _predict(state, dR, dt):
    """state (+) body-frame motion (dR, dt): pos += rot*dt, rot = rot*dR  (state layout of include/lio_b200.h)."""
    s = state.copy()
    Rw = s[3:12].reshape(3, 3)
    s[0:3] = s[0:3] + Rw @ dt
    s[3:12] = (Rw @ dR).reshape(9)
    return s


def test_sequence_with_map_growth_ate(pkg, orc, synth):
    """Config-1/3 shaped slice: scans along a street, IEKF update + map_incremental every scan, local-map cube moving.
    Two GPU contexts against one oracle:
      lock-step  - starts every scan from the ORACLE's state and inserts at the oracle's final state, so its inputs are
                   identical to the oracle's: per-iteration states within 1e-6 and the live map bit-identical;
      free-running - feeds its own estimates forward (its states differ from the oracle's by ~1e-15, which may flip the f32
                   rounding of an inserted world point): trajectory divergence must stay under 1 mm."""
    scene = synth.make_scene(half_extent=60.0)
    poses = synth.street_poses(12, scene, step=0.7)
    kw = dict(max_points=1 << 19, max_scan=1 << 16, cube_side_length=100.0, det_range=20.0)
    g, f = pkg.LioContext(0.5, **kw), pkg.LioContext(0.5, **kw)
    o = orc.OracleLio(0.5, nthreads=0, cube_side_length=100.0, det_range=20.0)
    R0, t0 = poses[0]
    first = synth.make_scan(scene, R0, t0, kind="vlp16", seed=100)
    world0 = (first.astype(np.float64) @ R0.T + t0).astype(np.float32)
    for c in (g, f, o):
        c.map_build(world0)
    fs = os_ = synth.make_state(R0, t0)
    fP = oP = synth.default_cov()
    Q = np.eye(23) * 1e-3
    traj_f, traj_o = [], []
    pred_rng = np.random.default_rng(77)
    for i, (R, t) in enumerate(poses[1:], 1):
        kind = "avia" if i % 2 else "vlp16"
        scan = synth.make_scan(scene, R, t, kind=kind, seed=100 + i, phase=0.37 * i)
        # IMU-like prediction (forward propagation is outside the hot path): compose the previous
        # estimate with the true relative motion plus a 2 cm / 0.2 deg error, inflate P
        Rp, tp = poses[i - 1]
        dR = Rp.T @ R @ synth.so3_exp(pred_rng.normal(0, 0.0035, 3))
        dt = Rp.T @ (t - tp) + pred_rng.normal(0, 0.02, 3)
        fs, os_ = _predict(fs, dR, dt), _predict(os_, dR, dt)
        fP, oP = fP + Q, oP + Q
        ob = o.fov_segment(os_[:3])
        assert np.array_equal(g.fov_segment(os_[:3]), ob)
        f.fov_segment(fs[:3])
        for c in (g, f, o):
            c.scan_upload(scan)
        gs, gP, gi = g.iekf_update(os_, oP, trace=True)
        fs, fP, _ = f.iekf_update(fs, fP)
        os_, oP, oi = o.iekf_update(os_, oP, trace=True)
        assert gi["iters"] == oi["iters"] and np.array_equal(gi["n_eff"], oi["n_eff"])
        assert np.abs(gi["states"] - oi["states"]).max() < STATE_TOL
        assert np.abs(gs - os_).max() < STATE_TOL and np.abs(fs - os_).max() < 1e-4
        assert g.map_incremental(os_, True) == o.map_incremental(os_, True)
        f.map_incremental(fs, True)
        traj_f.append(fs[:3].copy())
        traj_o.append(os_[:3].copy())
        if i % 4 == 0 or i == len(poses) - 1:
            gx, gid = g.map_export()
            ox, oid = o.map_export()
            assert np.array_equal(gid, oid) and np.array_equal(gx, ox)
            assert abs(f.map_size() - o.map_size()) <= 0.001 * o.map_size()
    ate = np.sqrt(np.mean(np.sum((np.array(traj_f) - np.array(traj_o)) ** 2, 1)))
    assert ate < 1e-3
    truth = np.array([t for _, t in poses[1:]])
    assert np.sqrt(np.mean(np.sum((np.array(traj_f) - truth) ** 2, 1))) < 0.08  # tracks ground truth too
    g.close()
    f.close()


def test_edge_cases_and_errors(pkg, synth, small_world):
    g = pkg.LioContext(0.5, max_points=4096, max_scan=256)
    st = small_world["truth"]
    with pytest.raises(pkg.LioError):  # no scan yet
        g.h_share(st)
    g.scan_upload(small_world["scan"][:100])
    with pytest.raises(pkg.LioError):  # cached neighbours requested before any search
        g.h_share(st, redo_knn=False)
    H, z, n_eff, res = g.h_share(st, True)  # empty map: valid == false, all-zero normal equations
    assert n_eff == 0 and not H.any() and not z.any() and res == 0
    s2, P2, info = g.iekf_update(st, synth.default_cov(), trace=True)
    assert np.array_equal(s2, st) and info["iters"] == 4 and (info["n_eff"] == 0).all()
    with pytest.raises(pkg.LioError):  # scan larger than max_scan
        g.scan_upload(small_world["scan"][:1000])
    ids, d2, pts, cnt = g.map_knn(np.zeros((3, 3), np.float32), 5, 5.0)
    assert (cnt == 0).all() and (ids == 0xFFFFFFFF).all() and np.isinf(d2).all()
    with pytest.raises(pkg.LioError):
        g.map_knn(np.zeros((3, 3), np.float32), 6, 5.0)
    assert g.map_add(np.zeros((0, 3), np.float32)) == 0
    # capacity exhaustion is reported, not silently dropped
    with pytest.raises(pkg.LioError):
        g.map_build(np.random.default_rng(0).uniform(-500, 500, (20000, 3)).astype(np.float32))
    g.close()
    with pytest.raises(pkg.LioError):
        pkg.LioContext(0.5, max_points=0, max_scan=16)


def test_full_size_properties(pkg, synth):
    """BASELINE config-2 sizes (131,072-point scan, 2M-point map): size-independent properties (the oracle comparison at
    this size is test_parity_at_baseline_config): kNN on a sample vs brute force, H^T H symmetric PSD and equal to the row
    products, determinism, idempotence."""
    rng = np.random.default_rng(11)
    M, N = 2_000_000, 131_072
    # planar-ish synthetic map: ground + two walls, one point per 0.5 m cell, jittered
    gx, gy = np.meshgrid(np.arange(-250, 250, 0.5), np.arange(-400, 400, 0.5), indexing="ij")
    ground = np.stack([gx.ravel(), gy.ravel(), np.zeros(gx.size)], 1)
    wy, wz = np.meshgrid(np.arange(-400, 400, 0.5), np.arange(0, 62.5, 0.5), indexing="ij")
    wall = np.stack([np.full(wy.size, 30.0), wy.ravel(), wz.ravel()], 1)
    m = np.concatenate([ground, wall, wall * [-1, 1, 1]])[:M]
    m = (m + rng.uniform(-0.2, 0.2, m.shape) * [1, 1, 0.05]).astype(np.float32)
    assert len(m) == M
    g = pkg.LioContext(0.5, max_points=M + (1 << 18), max_scan=N)
    g.map_build(m)
    assert g.map_size() == M
    az = rng.uniform(0, 2 * np.pi, N)
    r = rng.uniform(2, 100, N)
    scan = np.stack([r * np.cos(az), r * np.sin(az), np.full(N, -1.8)], 1)
    hit_wall = np.abs(scan[:, 0]) > 30
    scan[hit_wall, 0] = np.sign(scan[hit_wall, 0]) * 30
    scan[hit_wall, 2] = rng.uniform(-1.8, 40, hit_wall.sum())
    scan = (scan + rng.normal(0, 0.02, scan.shape)).astype(np.float32)
    st = synth.make_state(np.eye(3), [0, 0, 1.8])
    g.scan_upload(scan)
    H1, z1, n1, r1 = g.h_share(st, True)
    H2, z2, n2, r2 = g.h_share(st, True)
    assert n1 == n2 > 0.6 * N and np.array_equal(H1, H2) and np.array_equal(z1, z2)  # deterministic reduction
    assert np.array_equal(H1, H1.T) and np.linalg.eigvalsh(H1[:6, :6]).min() > 0
    sel, rows, z, n3 = g.h_share_rows(st, False)
    assert n3 == n1
    assert np.allclose(rows.T @ rows, H1, rtol=1e-9) and np.allclose(rows.T @ z, z1, rtol=1e-9, atol=1e-9)
    out = g.scan_get()
    idx = rng.integers(0, N, 64)
    for i in idx:
        d = m - out["world"][i]
        d2 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
        order = np.lexsort((np.arange(M), d2))[:5]
        order = order[d2[order] <= 5.0]
        assert out["nn_cnt"][i] == len(order)
        assert np.array_equal(out["nn_ids"][i, : len(order)], order.astype(np.uint32))
    # sortedness of every neighbour row
    ids, d2, _, cnt = g.map_knn(out["world"][:20000], 5, 5.0)
    assert (np.diff(d2, axis=1) >= 0).all()
    assert np.array_equal(ids, out["nn_ids"][:20000])
    # map_incremental twice: the second replay adds nothing that survives (idempotence of downsample-on-insert)
    g.map_incremental(st, True)
    s1 = g.map_size()
    g.h_share(st, True)
    a, b = g.map_incremental(st, True)
    assert g.map_size() - s1 <= b  # only no-downsample points may accumulate
    g.close()


def test_parity_at_baseline_config(pkg, orc, synth):
    """BASELINE configs[1] at full size (131,072-point OS1-128-shaped scan, 2M-point map), GPU vs the oracle: kNN index
    sets and plane masks bit-exact after a search and after a cached-neighbour call, H^T H / H^T z to 1e-5, n_eff and the state
    of every iteration (1e-6), and the live map after map_incremental as a set.  Same function bench.py puts in its line."""
    import bench

    w = bench.make_iekf_workload(synth, bench.N_SCAN, bench.N_MAP, 2)
    g = pkg.LioContext(0.5, max_points=bench.N_MAP + 8 * bench.N_SCAN, max_scan=bench.N_SCAN, max_cells=bench.N_MAP // 2)
    o = orc.OracleLio(0.5, nthreads=0)
    g.map_build(w["map"])
    o.map_build(w["map"])
    P0 = synth.default_cov()
    for j in range(2):  # the second scan runs against the map the first one grew
        par = bench.parity_at_config(g, o, w["scans"][j], w["starts"][j], w["truths"][j], P0)
        assert par["all"], par
        assert par["n_eff_first_call"] > 0.5 * bench.N_SCAN and par["live_map_points"] > bench.N_MAP
    g.close()
    o.close()
