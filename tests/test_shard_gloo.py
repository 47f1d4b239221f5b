"""Multi-rank host logic of the loop-ICP batch on CPU: world_size-2 gloo.  Each rank takes its contiguous
slice of the pair list, "verifies" it with the CPU oracle (the checker stands in for the GPU here: no GPU in this tier of
tests), and the results are gathered on the host exactly as loop_bench.loop_section does.  The union must equal the
single-rank answer, in pair order, with no pair lost or duplicated."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lb = importlib.import_module("bench_loop")
    synth = importlib.import_module("fast-lio-sam_gps_b200.synth")
    from oracle import orc

    w = lb.make_loop_workload(synth, n_pairs=5, n_pass=8, n_az=120, submap_range=2)
    idx = lb.shard_pair_indices(len(w["pairs"]), rank, world)
    res, _ = lb.cpu_loop(orc, w, [w["pairs"][i] for i in idx], 0.3, 1.5, 52.5, 1)
    mine = [(i, r["iterations"], bool(r["valid"]), float(r["score"]), r["T"].tobytes()) for i, r in zip(idx, res)]
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    dist.barrier()
    if rank == 0:
        q.put(sorted(x for part in gathered for x in part))  # back in pair order, as loop_section does
    dist.destroy_process_group()


def test_shard_pairs_partition():
    lb = importlib.import_module("bench_loop")
    for n in (0, 1, 5, 256, 257):
        for world in (1, 2, 3, 8):
            cuts = [lb.shard_pairs(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1


def test_shard_pair_indices_partition():
    """Two blocks per rank, dealt there and back: a partition of the list, sizes within 2, and the positions a rank owns
    average to the middle of the list (what cancels a linear cost trend)."""
    lb = importlib.import_module("bench_loop")
    for n in (0, 1, 5, 256, 257):
        for world in (1, 2, 3, 8):
            parts = [lb.shard_pair_indices(n, r, world) for r in range(world)]
            assert sorted(i for p in parts for i in p) == list(range(n))
            sizes = [len(p) for p in parts]
            assert max(sizes) - min(sizes) <= 2
            if n % (2 * world) == 0 and n:
                assert all(abs(np.mean(p) - (n - 1) / 2) < 1e-9 for p in parts)
    assert lb.shard_pair_indices(256, 0, 8) == list(range(0, 16)) + list(range(240, 256))


@pytest.mark.timeout(300)
def test_two_rank_gather_matches_single_rank():
    import torch.multiprocessing as mp

    lb = importlib.import_module("bench_loop")
    synth = importlib.import_module("fast-lio-sam_gps_b200.synth")
    from oracle import orc

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=240)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    w = lb.make_loop_workload(synth, n_pairs=5, n_pass=8, n_az=120, submap_range=2)
    ref, _ = lb.cpu_loop(orc, w, w["pairs"], 0.3, 1.5, 52.5, 1)
    assert [g[0] for g in got] == list(range(5))
    for g, r in zip(got, ref):
        assert g[1] == r["iterations"] and g[2] == bool(r["valid"]) and g[3] == float(r["score"]) and g[4] == r["T"].tobytes()
