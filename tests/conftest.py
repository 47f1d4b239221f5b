"""pytest configuration: the `gpu` marker, import helpers and shared seeded fixtures.

`-m "not gpu"` runs here on CPU (oracle vs golden vectors / brute force, host logic, C-ABI symbol checks);
`-m gpu` runs on a B200 and compares the CUDA path, called through the C-ABI, with the oracle.
"""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def pytest_collection_modifyitems(config, items):
    # a GPU test must never silently pass on a box without a device
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module("fast-lio-sam_gps_b200")


@pytest.fixture(scope="session")
def synth():
    return importlib.import_module("fast-lio-sam_gps_b200.synth")


@pytest.fixture(scope="session")
def orc():
    from oracle import orc as _orc

    _orc.build()
    return _orc


@pytest.fixture(scope="session")
def small_world(synth):
    """60 m scene, ~120k-point map, one VLP-16 scan with its true and perturbed state."""
    scene = synth.make_scene(half_extent=60.0)
    map_pts = synth.sample_map(scene, spacing=0.5)
    R = synth.rot_from_rpy(0.01, 0.02, 0.3)
    t = np.array([1.0, 0.5, 1.8])
    scan = synth.make_scan(scene, R, t, kind="vlp16")
    truth = synth.make_state(R, t)
    return dict(scene=scene, map=map_pts, scan=scan, truth=truth, start=synth.perturb_state(truth))


@pytest.fixture(scope="session")
def planted(synth):
    """Ground-plane map and a scan of it seen from 1.2 m; at `off` the sensor is 15 cm too high, so the weight test
    s = 1 - 0.9|pd2|/sqrt(range) > 0.9 fails for the points with sqrt(range) < 1.35 and passes for the far ones."""
    g = np.arange(-12, 12, 0.5, dtype=np.float32) + 0.25
    gx, gy = np.meshgrid(g, g)
    m = np.stack([gx.ravel(), gy.ravel(), np.full(gx.size, 5.0, np.float32)], 1)  # (esti_plane cannot express a plane through the origin)
    rng = np.random.default_rng(7)
    xy = rng.uniform(-8, 8, (3000, 2)).astype(np.float32)
    scan = np.concatenate([xy, np.full((len(xy), 1), -1.2, np.float32)], 1)  # body frame = world shifted by the height
    truth = synth.make_state(np.eye(3), np.array([0.0, 0.0, 6.2]))
    off = synth.make_state(np.eye(3), np.array([0.0, 0.0, 6.35]))
    return m, scan, truth, off

