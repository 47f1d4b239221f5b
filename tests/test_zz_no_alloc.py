"""lio_hth_reduce_static (include/lio_b200.h): the allocation-free variant of the H^T H / H^T z reduction, whose scratch is
a static __device__ array.  Checked against the existing entry point lio_h_share on the same scan, on calls in a row (the
array keeps the previous call's contents).  Collected last on purpose.
"""
import numpy as np
import pytest

REL = 1e-11  # FP64 sums in a different order; same bar as the H^T H comparison in test_gpu_lio.py


def rel_err(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.gpu
def test_static_reduce_matches_h_share(pkg, small_world):
    import torch

    w = small_world
    dev = torch.device("cuda:0")
    g = pkg.LioContext(0.5, max_points=1 << 19, max_scan=1 << 16)
    g.map_build(w["map"])
    g.scan_upload(w["scan"])
    # the caller owns inputs and outputs; outputs start as garbage so a call that wrote nothing cannot pass
    hth = torch.full((12, 12), float("nan"), dtype=torch.float64, device=dev)
    htz = torch.full((12,), float("nan"), dtype=torch.float64, device=dev)
    n_full = len(w["scan"])
    assert n_full > 3 * 1024  # several blocks of partial sums
    # a long call, a short one (fewer blocks than the scratch last held), the long one again, then an empty one
    for call, (state, n_rows) in enumerate(((w["start"], n_full), (w["truth"], 700), (w["start"], n_full), (w["truth"], 0))):
        ref_H, ref_z, n_eff, _ = g.h_share(state, True)
        sel, rows, z, n_r = g.h_share_rows(state, True)
        assert n_r == n_eff > 0
        if n_rows < n_full:  # the existing entry point on a prefix = the plain sum over the prefix
            rows, z = rows[:n_rows].copy(), z[:n_rows].copy()
            ref_H, ref_z = rows.T @ rows, rows.T @ z
        rows_d = torch.from_numpy(rows).to(dev)
        z_d = torch.from_numpy(z).to(dev)
        torch.cuda.synchronize()
        before = (torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0])
        pkg.hth_reduce_static(rows_d.data_ptr(), z_d.data_ptr(), n_rows, hth.data_ptr(), htz.data_ptr(),
                              torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before[0]
        if call > 0:  # (the first launch may make the driver load the module that holds the kernel and the array)
            assert torch.cuda.mem_get_info()[0] == before[1]
        got_H, got_z = hth.cpu().numpy(), htz.cpu().numpy()
        if n_rows == 0:
            assert (got_H == 0).all() and (got_z == 0).all()
            continue
        assert np.array_equal(got_H, got_H.T)
        assert rel_err(got_H, ref_H) < REL and rel_err(got_z, ref_z) < 1e-9
        hth.fill_(float("nan"))
        htz.fill_(float("nan"))
    g.close()


@pytest.mark.gpu
def test_static_reduce_rejects_oversize(pkg):
    import torch

    out = torch.zeros(156, dtype=torch.float64, device="cuda:0")
    with pytest.raises(pkg.LioError, match="LIO_STATIC_REDUCE_MAX_ROWS"):
        pkg.hth_reduce_static(out.data_ptr(), out.data_ptr(), pkg.STATIC_REDUCE_MAX_ROWS + 1, out.data_ptr(), out.data_ptr() + 144 * 8)
    with pytest.raises(pkg.LioError, match="null"):
        pkg.hth_reduce_static(None, None, 8, out.data_ptr(), out.data_ptr() + 144 * 8)


@pytest.mark.gpu
def test_static_reduce_at_the_limit(pkg):
    """LIO_STATIC_REDUCE_MAX_ROWS rows (every block of the scratch in use) against a float64 torch product."""
    import torch

    n = pkg.STATIC_REDUCE_MAX_ROWS
    gen = torch.Generator(device="cuda:0").manual_seed(5)
    rows = torch.randn((n, 12), dtype=torch.float64, device="cuda:0", generator=gen)
    z = torch.randn((n,), dtype=torch.float64, device="cuda:0", generator=gen)
    hth = torch.empty((12, 12), dtype=torch.float64, device="cuda:0")
    htz = torch.empty((12,), dtype=torch.float64, device="cuda:0")
    for _ in range(2):
        pkg.hth_reduce_static(rows.data_ptr(), z.data_ptr(), n, hth.data_ptr(), htz.data_ptr(), None)
        torch.cuda.synchronize()
        assert rel_err(hth.cpu().numpy(), (rows.T @ rows).cpu().numpy()) < REL
        assert rel_err(htz.cpu().numpy(), (rows.T @ z).cpu().numpy()) < 1e-9  # H^T z is a sum of mixed signs near zero
        hth.zero_()
