#!/usr/bin/env python
"""Headline benchmark: IEKF scans/s on Ouster OS1-128-shaped scans (131,072 pts) against a 2M-point map
(BASELINE.json configs[1]), and `--workload loop` for the batched loop-closure ICP (configs[3]).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload iekf|loop]

One "step" = one scan through the hot path: esekfom's update_iterated_dyn_share_modified (up to
max_iteration+1 = 4 calls of the GPU h_share_model: body->world, exact k=5 NN, plane fit, residual,
Jacobian, FP64 H^T H / H^T z) followed by map_incremental (classification + downsample-on-insert).
`value` times it with the scan already resident in HBM; `e2e` times the same call sequence through the
C-ABI with HOST (pinned) scan buffers, the H2D copy and the D2H read of H^T H / H^T z inside the region.
N > 1: the IEKF does not shard (iterations are sequential) -> replicas only, one independent sequence per
GPU, no collective; scaling is weak.
`--impl reference` times the CPU oracle (oracle/, the restatement of the reference's OpenMP path; the
reference itself cannot be built here) on the host cores on the same workload.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_SCAN = 131072  # 128 x 1024
N_MAP = 2_000_000
ALGO_BYTES_PER_POINT = 96  # 16 B body point + 5 x 16 B neighbours, per query point per call
MAX_ITER = 3


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="iekf", choices=["iekf", "loop", "knn_sweep", "map_growth", "sequence"])
    ap.add_argument("--scans", type=int, default=4, help="distinct synthetic scans cycled through")
    ap.add_argument("--n-scan", type=int, default=N_SCAN)
    ap.add_argument("--n-map", type=int, default=N_MAP)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pairs", type=int, default=256, help="loop workload: candidate pairs in the batch")
    ap.add_argument("--submap-range", type=int, default=12, help="loop workload: R (2R+1 keyframes per side)")
    ap.add_argument("--n-pass", type=int, default=100, help="loop workload: keyframes per pass of the out-and-back drive")
    ap.add_argument("--sweep-map", type=int, default=20_000_000, help="knn_sweep workload: map points")
    ap.add_argument("--sweep-queries", type=int, default=100_000, help="knn_sweep workload: queries per call")
    ap.add_argument("--sweep-big-queries", type=int, default=1_000_000, help="knn_sweep workload: queries per call of the extra rows at filter_size_map 0.5 (0: skip)")
    ap.add_argument("--seq-scans", type=int, default=100, help="sequence workload: VLP-16 scans (configs[0] names 1000)")
    ap.add_argument("--map-scans", type=int, default=200, help="map_growth workload: Avia-shaped scans along the drive")
    ap.add_argument("--map-check-every", type=int, default=50, help="map_growth workload: compare the live set with the oracle every N scans")
    ap.add_argument("--no-loop", action="store_true", help="iekf workload: skip the loop_icp section of the JSON line")
    ap.add_argument("--no-parity", action="store_true", help="iekf workload: skip the oracle-vs-GPU comparison at the full config")
    ap.add_argument("--no-configs", action="store_true", help="iekf workload: skip the short runs of configs[0] / [2] / [4] folded into the JSON line")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ------------------------------------------------------------------------------------------- workload
def make_iekf_workload(synth, n_scan, n_map, n_scans, rank=0):
    """Seeded street scene big enough for an n_map-point map at one point per 0.5 m cell; OS1-128 scans."""
    half = 60.0
    while True:
        scene = synth.make_scene(half_extent=half, seed=synth.SEED + 1000 * rank)
        # surface area estimate: ground + roofs ~ (2*half)^2, walls ~ 4 * w * h per block
        nb = len(scene["boxes"])
        est = ((2 * half) ** 2 + nb * 4 * 8.0 * 14.0) / 0.25
        if est >= n_map * 1.05 or half > 600:
            break
        half += 20.0
    m = synth.sample_map(scene, spacing=0.5, seed=synth.SEED + 1 + 1000 * rank)
    while len(m) < n_map:  # rare: top up with a second, offset sampling
        extra = synth.sample_map(scene, spacing=0.5, seed=synth.SEED + 7 + len(m))
        m = np.concatenate([m, extra[: n_map - len(m)]])
    m = np.ascontiguousarray(m[:n_map])
    scans, truths, starts = [], [], []
    for i in range(n_scans):
        R = synth.rot_from_rpy(0.01 * i, -0.01, 0.2 + 0.3 * i)
        t = np.array([-0.25 * half + 3.0 * i, 0.4 * i, 1.9])
        s = synth.make_scan(scene, R, t, kind="os1_128", n_points=n_scan, seed=synth.SEED + 50 + i, max_range=min(120.0, half))
        if len(s) < n_scan:  # sky-heavy pose: repeat returns with fresh noise so the scan has exactly n_scan points
            rng = np.random.default_rng(i)
            idx = rng.integers(0, len(s), n_scan - len(s))
            s = np.concatenate([s, s[idx] + rng.normal(0, 0.01, (len(idx), 3)).astype(np.float32)])
        scans.append(np.ascontiguousarray(s[:n_scan]))
        st = synth.make_state(R, t)
        truths.append(st)
        starts.append(synth.perturb_state(st, seed=synth.SEED + 90 + i))
    return dict(scene=scene, map=m, scans=scans, truths=truths, starts=starts, half=half)


# ------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi poller (one process, one GPU).  Rows are stamped with the host clock so that stop(t0, t1) can keep the
    samples that fell INSIDE the measured window."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index, period_ms=100):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", str(period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def wait_first(self, timeout=3.0):
        t_end = time.perf_counter() + timeout
        while self.proc and not self.rows and time.perf_counter() < t_end:
            time.sleep(0.02)

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        rows = [r for ts, r in self.rows if len(r) >= 6 and (t0 is None or t0 <= ts <= t1)]
        inside = len(rows)
        if not rows:  # window shorter than the polling period: fall back to the sample nearest the window
            rows = [r for _, r in self.rows if len(r) >= 6][-1:]
        sm = [float(r[0]) for r in rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[j] for r in rows for j in range(4) if r[2 + j].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": inside}


# ------------------------------------------------------------------------------------------- parity at the BASELINE config
REL_TOL, STATE_TOL = 1e-5, 1e-6  # north_star: H^T H / H^T z to 1e-5 relative, per-iteration states to 1e-6 m / rad


def parity_at_config(g, o, scan, start, truth, P0, max_iter=MAX_ITER):
    """GPU path (C-ABI) against the CPU oracle on ONE full-size scan of the workload, both starting from the same map:
    kNN index sets and plane-validity masks bit-exact after a search call and after a cached-neighbour call, H^T H / H^T z
    to 1e-5 relative, n_eff and the state of every IEKF iteration (1e-6), and the live map after map_incremental as a
    set of (id, xyz).  Mutates both maps identically (one scan inserted).  Returns a dict of booleans + the worst errors."""
    rel = lambda a, b: float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
    out = {}
    g.scan_upload(scan)
    o.scan_upload(scan)
    worst_h = 0.0
    knn_ok = mask_ok = neff_ok = hth_ok = True
    prev = None
    for redo, st in ((True, start), (False, truth)):
        gH, gz, gn, _ = g.h_share(st, redo)
        oH, oz, on, _ = o.h_share(st, redo)
        go, oo = g.scan_get(), o.scan_get()
        knn_ok &= bool(np.array_equal(go["nn_cnt"], oo["nn_cnt"]) and np.array_equal(go["nn_ids"], oo["nn_ids"]))
        mask_ok &= bool(np.array_equal(go["sel"], oo["sel"]))
        if prev is not None:
            mask_ok &= bool((go["sel"] <= prev).all())  # point_selected_surf only shrinks between searches
        prev = go["sel"].copy()
        neff_ok &= gn == on
        e = max(rel(gH, oH), rel(gz, oz))
        worst_h = max(worst_h, e)
        hth_ok &= e < REL_TOL
    out.update(knn_sets=knn_ok, plane_masks=mask_ok, hth_htz_rel_1e5=hth_ok, hth_htz_worst_rel=worst_h, n_eff_first_call=int(gn))
    g.scan_upload(scan)
    o.scan_upload(scan)
    gs, gP, gi = g.iekf_update(start, P0, max_iter=max_iter, trace=True)
    os_, oP, oi = o.iekf_update(start, P0, max_iter=max_iter, trace=True)
    # map_incremental is compared as a function of (state, cached neighbours, map): both sides get the ORACLE's updated state.
    # (The two states agree to ~1e-15, but fed separately that last bit can flip the f32 rounding of an inserted point, after
    # which "live set equality" would be testing the chaos of rounding, not the insert semantics.)
    n_lists_g = g.map_incremental(os_, True)
    n_lists_o = o.map_incremental(os_, True)
    same_trace = gi["iters"] == oi["iters"] and np.array_equal(gi["n_eff"], oi["n_eff"]) and np.array_equal(gi["knn"], oi["knn"])
    st_err = float(np.abs(gi["states"] - oi["states"]).max()) if gi["iters"] == oi["iters"] else float("inf")
    out.update(n_eff_per_iteration=bool(same_trace) and bool(neff_ok), states_1e6=st_err < STATE_TOL, state_worst_abs=st_err,
               cov_worst_abs=float(np.abs(gP - oP).max()), iterations=int(gi["iters"]))
    gx, gid = g.map_export()
    ox, oid = o.map_export()
    ga, oa = np.argsort(gid, kind="stable"), np.argsort(oid, kind="stable")
    out["live_map_set"] = bool(len(gid) == len(oid) and np.array_equal(gid[ga], oid[oa]) and np.array_equal(gx[ga], ox[oa]))
    out["live_map_points"] = int(len(gid))
    out["map_incremental_lists"] = [int(x) for x in n_lists_g]
    out["live_map_set"] = bool(out["live_map_set"] and tuple(n_lists_g) == tuple(n_lists_o))
    out["all"] = bool(all(out[k] for k in ("knn_sets", "plane_masks", "hth_htz_rel_1e5", "n_eff_per_iteration", "states_1e6", "live_map_set")))
    return out


# ------------------------------------------------------------------------------------------- CPU legs
def cpu_iekf(orc, synth, w, steps, warmup, with_map_update=True):
    """The oracle (CPU restatement of the reference's OpenMP h_share_model + ikd-Tree + esekfom) on the same scans."""
    cores = orc.max_threads()
    o = orc.OracleLio(0.5, nthreads=cores)
    o.map_build(w["map"])
    P0 = synth.default_cov()
    calls = 0
    n_eff = 0
    for i in range(warmup + steps):
        if i == warmup:
            t0 = time.perf_counter()
            calls = n_eff = 0
        j = i % len(w["scans"])
        o.scan_upload(w["scans"][j])
        s, P, info = o.iekf_update(w["starts"][j], P0, max_iter=MAX_ITER, trace=True)
        calls += info["iters"]
        n_eff += int(info["n_eff"].sum())
        if with_map_update:
            o.map_incremental(s, True)
    dt = time.perf_counter() - t0
    o.close()
    return dict(scans_per_s=steps / dt, ms_per_step=1e3 * dt / steps, cores=cores, calls_per_scan=calls / steps, corr_per_s=n_eff / dt)


def run_reference(args, synth):
    from oracle import orc

    rank, _, world = dist_env()
    if rank != 0:
        return
    orc.build()
    w = make_iekf_workload(synth, args.n_scan, args.n_map, args.scans)
    r = cpu_iekf(orc, synth, w, args.steps, max(1, min(args.warmup, 2)))
    line = {
        "impl": "reference", "metric": "IEKF scans/s", "value": r["scans_per_s"], "unit": "scans/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32 search/plane + f64 Jacobian/reduction", "data": "synthetic",
        "config": workload_config(args, w, r["calls_per_scan"]),
        "cpu_baseline": {"value": r["scans_per_s"], "unit": "scans/s", "cores": r["cores"], "kind": "port",
                         "sample": f"{args.steps} full scans ({args.n_scan} pts x {r['calls_per_scan']:.2f} model calls + map_incremental) "
                                   f"on the {args.n_map}-pt map; reference source absent -> oracle port, OpenMP over all host cores"},
        "e2e": {"value": r["scans_per_s"], "unit": "scans/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "corr_per_s": r["corr_per_s"], "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args, w, calls):
    return {"workload": "configs[1]: OS1-128-shaped synthetic scans, IEKF h_share_model + map_incremental against a 2M-point map",
            "points_per_scan": args.n_scan, "map_points": args.n_map, "k": 5, "max_iteration": MAX_ITER,
            "model_calls_per_scan": round(float(calls), 3), "filter_size_map": 0.5, "distinct_scans": len(w["scans"]),
            "state_perturbation": "5 cm / 0.5 deg", "l2_policy": "map (table+buckets > 126 MB L2 not guaranteed: see DESIGN.md); "
            "a 256 MB buffer is written between timed steps to flush L2", "parallelism": "replicas (one sequence per GPU)"}


# ------------------------------------------------------------------------------------------- GPU leg
def run_ours(args, synth):
    import torch
    import torch.distributed as dist

    pkg = importlib.import_module("fast-lio-sam_gps_b200")
    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"  # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if not os.path.exists(pkg.LIB_PATH):
        pkg.build_library()

    w = make_iekf_workload(synth, args.n_scan, args.n_map, args.scans, rank=rank)
    stream = torch.cuda.Stream()
    g = pkg.LioContext(0.5, max_points=args.n_map + 8 * args.n_scan, max_scan=args.n_scan, max_cells=max(1 << 16, args.n_map // 2), device=local_rank,
                       stream=stream.cuda_stream)  # ~0.38 occupied 1 m cells per map point in this scene -> table load 0.37
    g.map_build(w["map"])
    P0 = synth.default_cov()
    ns = len(w["scans"])
    # device-resident scans (float4) for the `value` leg, pinned host copies (xyz) for the `e2e` leg
    dev_scans, pin_scans = [], []
    for s in w["scans"]:
        s4 = np.concatenate([s, np.zeros((len(s), 1), np.float32)], 1)
        dev_scans.append(torch.from_numpy(s4).cuda())
        pin_scans.append(torch.from_numpy(s.copy()).pin_memory())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(i, host):
        j = i % ns
        if host:
            g.scan_upload_ptr(pin_scans[j].data_ptr(), args.n_scan)  # consumes the copy prefetched during the previous step
            # the NEXT step's scan starts its H2D copy now, on the copy stream, under this step's kernels (lio_scan_prefetch):
            # one 1.57 MB copy per step stays inside the timed region, it just no longer sits in front of the first kernel
            g.scan_prefetch_ptr(pin_scans[(j + 1) % ns].data_ptr(), args.n_scan)
        else:
            g.scan_set_device(dev_scans[j].data_ptr(), args.n_scan)
        # update_iterated_dyn_share_modified + map_incremental in one enqueue (lio_iekf_update_map): the device-resident loop
        s, P, info = g.iekf_update_map(w["starts"][j], P0, True, max_iter=MAX_ITER, trace=True)
        return info

    def timed(host, steps, warmup):
        for i in range(warmup):
            step(i, host)
        g.sync()
        g.timing(True)
        g.timing_reset()
        l0 = g.launch_count()
        evs = []
        calls = n_eff = 0
        barrier()
        for i in range(steps):
            with torch.cuda.stream(stream):
                flush.fill_(i & 255)  # L2 flush between timed steps (outside the event pair)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                info = step(warmup + i, host)
                b.record(stream)
            evs.append((a, b))
            calls += info["iters"]
            n_eff += int(info["n_eff"].sum())
        g.sync()
        barrier()
        total_ms = sum(a.elapsed_time(b) for a, b in evs)
        launches = g.launch_count() - l0
        kt = {k: g.timing_get(k) for k in ("sort", "knn", "plane_reduce", "insert", "classify")}
        g.timing(False)
        return total_ms, calls, n_eff, launches, kt

    def wall_leg(n_scans):
        """Cross-check of the event-timed legs: n_scans scans back to back through the host-buffer path (upload + prefetch +
        lio_iekf_update_map), no L2 flush, no per-kernel events, host wall clock around the loop.  Long enough (about a second)
        for nvidia-smi to see the clocks UNDER this workload: rank 0 polls them here, not inside the 8 ms event-timed legs (eight
        pollers inside the timed leg cost the round-1 replicas run 9 % at N = 8)."""
        for i in range(3):
            step(i, True)
        g.sync()
        clk = ClockSampler(local_rank) if rank == 0 else None
        if clk:
            clk.wait_first()
        barrier()
        t0 = time.perf_counter()
        for i in range(n_scans):
            step(3 + i, True)
        g.sync()
        t1 = time.perf_counter()
        barrier()
        return t1 - t0, (clk.stop(t0, t1) if clk else None)

    parity = None
    if rank == 0 and world == 1 and not args.no_parity:  # oracle vs GPU on one full-size scan of THIS workload, before the map is touched
        from oracle import orc

        orc.build()
        o = orc.OracleLio(0.5, nthreads=orc.max_threads())
        o.map_build(w["map"])
        parity = parity_at_config(g, o, w["scans"][0], w["starts"][0], w["truths"][0], P0)
        o.close()

    st0 = g.knn_stats_detail()
    dev_ms, calls, n_eff, launches, kt = timed(False, args.steps, args.warmup)
    st1 = g.knn_stats_detail()
    knn_calls = max(1, kt["knn"][1]) * (args.warmup + args.steps) / max(1, args.steps)  # the counters also saw the warm-up
    knn_deferred = {k: (st1[k] - st0[k]) / knn_calls / args.n_scan for k in st0}  # fraction of a scan's queries, per search
    e2e_ms, e_calls, _, _, _ = timed(True, args.steps, max(3, args.warmup))
    wall_scans = max(200, min(4000, int(1.2e3 / max(e2e_ms / args.steps, 1e-3))))  # ~1.2 s of scans
    wall_s, clocks = wall_leg(wall_scans)
    t = torch.tensor([dev_ms, e2e_ms, wall_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)  # max over ranks of the device-timed totals
    dev_ms, e2e_ms, wall_s = t.tolist()
    total_scans = args.steps * world
    value = total_scans / (dev_ms * 1e-3)
    e2e = total_scans / (e2e_ms * 1e-3)

    loop_out = None
    if not args.no_loop:  # configs[3] in the same run: pairs sharded over the ranks, host gather only
        loop_mod = importlib.import_module("fast-lio-sam_gps_b200.loop")
        loop_bench = importlib.import_module("bench_loop")
        loop_out = loop_bench.loop_section(synth, loop_mod, rank, local_rank, world, args.pairs, 5, 1, submap_range=args.submap_range,
                                           n_pass=args.n_pass, cpu_pairs=0 if (args.no_cpu_baseline or world > 1) else 16, stream=stream)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        traffic = None  # DRAM bytes per search from the committed ncu --set full capture (profiles/traffic.json)
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["iekf_knn"]
            traffic = {"bytes_per_launch": tr["bytes_per_launch"], "source": tr["source"]} if tr and args.n_scan == N_SCAN and args.n_map == N_MAP else None
        except Exception:
            pass
        knn_ms, knn_n = kt["knn"]
        algo_bytes = ALGO_BYTES_PER_POINT * args.n_scan
        achieved = algo_bytes / (knn_ms * 1e-3) / 1e9 if knn_ms > 0 else 0.0
        line = {
            "metric": "IEKF scans/s", "value": value, "unit": "scans/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 search/plane + f64 Jacobian/reduction", "data": "synthetic",
            "config": workload_config(args, w, calls / args.steps),
            "corr_per_s": n_eff * world / (dev_ms * 1e-3),
            # H2D: the scan (xyz f32) + the IekfIO input block (state, P, limit, R); D2H: state, P and the per-call traces the
            # device loop writes into the mapped result block
            "e2e": {"value": e2e, "unit": "scans/s", "h2d": "pinned host scan -> lio_scan_upload; the next step's scan is prefetched on a copy "
                    "stream (lio_scan_prefetch) inside the current step's timed region", "h2d_bytes_per_step": args.n_scan * 12 + (36 + 529 + 23 + 2) * 8,
                    "d2h_bytes_per_step": int(round((36 + 529) * 8 + e_calls / args.steps * (36 * 8 + 8) + 16))},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "k_knn_tile + k_knn_left (body->world + exact k=5 NN; Morton sort on the first call of a scan)", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic["bytes_per_launch"] if traffic else None,
                         "traffic_source": traffic["source"] if traffic else None,
                         "algorithmic_bytes_per_launch": algo_bytes, "kernel_ms": knn_ms, "launches_timed": knn_n,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst copy)" if peaks else "fallback 6650"},
            "kernel_ms": {k: {"mean_ms": v[0], "launches": v[1]} for k, v in kt.items()},
            "knn_deferred_fraction": knn_deferred,
            "clocks": dict(clocks or {}, sampled="rank 0, nvidia-smi -lms 100 during the e2e_wall leg (same workload, back to back)"),
            "e2e_wall": {"value": wall_scans * world / wall_s, "unit": "scans/s", "scans": wall_scans, "seconds": wall_s,
                         "timing": "time.perf_counter around back-to-back scans through the host-buffer path, no L2 flush, max over ranks"},
        }
        if parity is not None:
            line["parity_at_config"] = parity
        if not args.no_cpu_baseline and world == 1:
            from oracle import orc

            orc.build()
            r = cpu_iekf(orc, synth, w, steps=max(2, min(6, args.steps)), warmup=1)
            line["cpu_baseline"] = {"value": r["scans_per_s"], "unit": "scans/s", "cores": r["cores"], "kind": "port",
                                    "sample": f"{max(2, min(6, args.steps))} full scans of the same workload (oracle port, OpenMP, all host cores)",
                                    "ms_per_step": r["ms_per_step"]}
        if loop_out is not None:
            line["loop_icp"] = loop_out
    g.close()
    if rank == 0:
        if world == 1 and not args.no_configs:
            line["configs"] = compact_configs(args)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and not parity["all"]:
        raise SystemExit("bench.py: the GPU path differs from the oracle at the BASELINE config: " + json.dumps(parity))


def compact_configs(args):
    """Short runs of the other BASELINE.json configs, folded into the headline line so that the driver-run record carries all
    five (the full-size runs are the standalone workloads: --workload sequence / map_growth / knn_sweep / loop).  Each entry:
    value, cpu_baseline, parity flag, roofline fraction where the workload has one.  Bounded to about half a minute."""
    import copy

    out = {}
    t0 = time.perf_counter()
    a = copy.copy(args)
    a.seq_scans = 30
    try:
        r = importlib.import_module("bench_seq").run(a)
        out["configs[0] sequence"] = {"value": r["value"], "unit": r["unit"], "scans": r["config"]["scans"], "full_size": "--workload sequence --seq-scans 1000",
                                      "cpu_baseline": r["cpu_baseline"]["value"], "ate_difference_m": r["ate_difference_m"],
                                      "lockstep_position_difference_max_m": r["lockstep"]["position_difference_max_m"],
                                      "free_running_divergence_max_m": r["trajectory_divergence_max_m"],
                                      "parity": bool(r["ate_difference_m"] < r["bar_m"] and r["lockstep"]["position_difference_max_m"] < r["bar_m"])}
    except Exception as e:  # a short run must not take the headline line down with it
        out["configs[0] sequence"] = {"error": str(e)[:200]}
    a = copy.copy(args)
    a.map_scans, a.map_check_every = 41, 20
    try:
        r = importlib.import_module("bench_map").run(a)
        out["configs[2] map_growth"] = {"value": r["value"], "unit": r["unit"], "scans": r["config"]["scans"], "full_size": "--workload map_growth --map-scans 2001",
                                        "cpu_baseline": r["cpu_baseline"]["value"], "parity": bool(r["live_set_identical_to_oracle"]),
                                        "roofline_frac": r["roofline"]["frac"]}
    except Exception as e:
        out["configs[2] map_growth"] = {"error": str(e)[:200]}
    a = copy.copy(args)
    a.steps = 3
    try:
        r = importlib.import_module("bench_sweep").run(a, spacings=(0.5,))
        row = r["sweep"][0]
        big = {b["query_set"]: b for b in r["sweep_1M"]}
        out["configs[4] knn_sweep"] = {"value": row["queries_per_s"], "unit": "queries/s", "map_points": r["config"]["map_points"], "filter_size_map": 0.5,
                                       "full_size": "--workload knn_sweep (four densities)", "parity": bool(row["exact_on_sample"] and all(b["exact_on_sample"] for b in big.values())),
                                       "fraction_of_gather_probe_100k": row["fraction_of_gather_probe_16B"],
                                       "queries_per_s_1M_patch": big["patch"]["queries_per_s"], "fraction_of_gather_probe_1M_patch": big["patch"]["fraction_of_gather_probe_16B"],
                                       "queries_per_s_1M_spread_l2_flushed": big["spread"]["queries_per_s_l2_flushed"],
                                       "fraction_of_gather_probe_1M_spread": big["spread"]["fraction_of_gather_probe_16B"],
                                       "dram_traffic_1M_spread": big["spread"].get("dram_traffic"),
                                       "gather_probe_GBps_16B": r["gather_probe"]["GBps_16B_requests"]}
    except Exception as e:
        out["configs[4] knn_sweep"] = {"error": str(e)[:200]}
    out["seconds"] = time.perf_counter() - t0
    return out


def main():
    args = parse()
    synth = importlib.import_module("fast-lio-sam_gps_b200.synth")
    if args.workload == "loop":
        loop_bench = importlib.import_module("bench_loop")
        return loop_bench.main(args)
    if args.workload == "sequence":
        return importlib.import_module("bench_seq").main(args)
    if args.workload == "map_growth":
        return importlib.import_module("bench_map").main(args)
    if args.workload == "knn_sweep":
        return importlib.import_module("bench_sweep").main(args)
    if args.impl == "reference":
        return run_reference(args, synth)
    return run_ours(args, synth)


if __name__ == "__main__":
    main()
