"""pytest configuration: the `gpu` marker, in-tree native builds, shared workloads.

`-m "not gpu"` covers the oracle (against the committed golden fixtures generated from the reference itself), the
host-side logic, and that the C-ABI library loads and exports every symbol include/mins_lidar.h declares.
`-m gpu` tests are the parity tests proper: CUDA path (through the C ABI) vs the oracle on the same seeded inputs.
Nothing here reads /root/reference at run time.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def native_build():
    """Build (if stale) libmins_lidar.so, liboracle.so and — where the reference tree exists — oracle/_ref."""
    import __graft_entry__ as g
    g.build()


@pytest.fixture(scope="session")
def po():
    from oracle import pyoracle
    pyoracle.lib()
    return pyoracle


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def cuda_ok():
    # GPU tests must FAIL (not skip) on a box that should have a GPU; they are simply deselected with -m "not gpu".
    assert _has_cuda(), "this test needs a CUDA device; there is no CPU fallback in mins_b200"
    return True


_WL_CACHE = {}


@pytest.fixture(scope="session")
def workload():
    """cached synthetic workloads keyed by (shape, variant, k, downsample, map_points)"""
    from mins_b200 import synth
    from mins_b200.lidar import OptionsLidar

    def get(shape="G-small", variant="plain", k=5, downsample=False, map_points=None, seed=None):
        key = (shape, variant, k, downsample, map_points, seed)
        if key not in _WL_CACHE:
            op = OptionsLidar(raw_do_downsample=downsample, map_ngbr_num=k)
            _WL_CACHE[key] = synth.make_workload(shape, op=op, variant=variant, map_points=map_points, seed=seed)
        return _WL_CACHE[key]

    return get


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
