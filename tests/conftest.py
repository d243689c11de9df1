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

    def get(shape="G-small", variant="plain", k=5, downsample=False, map_points=None, seed=None, options="sim"):
        key = (shape, variant, k, downsample, map_points, seed, options)
        if key not in _WL_CACHE:
            if options == "kaist":  # mins/config/kaist/config_lidar.yaml: the map reaches map_decay_dist = 30 m
                from mins_b200.lidar import kaist_options
                op = kaist_options(raw_do_downsample=downsample, map_ngbr_num=k)
                _WL_CACHE[key] = synth.make_workload(shape, op=op, variant=variant, map_points=map_points, seed=seed,
                                                     map_radius=op.map_decay_dist)
            else:
                op = OptionsLidar(raw_do_downsample=downsample, map_ngbr_num=k)
                _WL_CACHE[key] = synth.make_workload(shape, op=op, variant=variant, map_points=map_points, seed=seed)
        return _WL_CACHE[key]

    return get


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def check_voxel_contract(src, out, leaf, po, post=None, slack=1.0):
    """The stated contract of the PCL VoxelGrid restatement: `out` holds one centroid per occupied
    voxel of `src`, in ascending voxel index (identical membership and order as the oracle), each within
    4 ulp * sqrt(n_voxel) of the oracle's centroid (fp32 accumulation order inside a voxel is not defined by PCL either).
    post(ref) -> (ref2, keep) optionally maps the oracle centroids through a later stage (re-anchor: decay filter + transform);
    slack scales the tolerance for that stage's own rounding. Returns the per-voxel point counts."""
    ref, vid = po.voxel_downsample(src, leaf)
    assert (np.diff(vid.astype(np.int64)) > 0).all()
    inv = np.float32(1.0) / np.float32(leaf)
    mn = np.floor(src[:, :3].min(0) * inv).astype(np.int64)
    dv = np.floor(src[:, :3].max(0) * inv).astype(np.int64) - mn + 1
    lin = ((np.floor(src[:, 0] * inv).astype(np.int64) - mn[0]) + (np.floor(src[:, 1] * inv).astype(np.int64) - mn[1]) * dv[0]
           + (np.floor(src[:, 2] * inv).astype(np.int64) - mn[2]) * dv[0] * dv[1])
    pos = np.searchsorted(vid, lin)
    assert (vid[pos] == lin).all()                                   # the oracle's voxel ids are the PCL linear indices
    counts = np.bincount(pos, minlength=len(vid))
    assert counts.sum() == len(src) and (counts > 0).all()
    tol = 4 * np.finfo(np.float32).eps * np.sqrt(counts)[:, None] * np.maximum(np.abs(ref).max(1), 1.0)[:, None]
    if post is not None:
        ref, keep = post(ref)
        tol = tol[keep]
        tol = slack * tol + np.finfo(np.float32).eps * np.maximum(np.abs(ref).max(1), 1.0)[:, None]
    assert len(out) == len(ref), (len(out), len(ref))                # same occupied voxels (membership) ...
    err = np.abs(out.astype(np.float64) - ref.astype(np.float64))
    assert (err <= tol).all(), float((err / tol).max())              # ... in the same order, centroids within the bound
    return counts
