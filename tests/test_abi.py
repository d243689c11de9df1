"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports exactly what include/mins_lidar.h declares,
struct layouts agree between the header (as compiled into the .so and the oracle) and the ctypes mirror, the library
refuses to run without a GPU (no CPU fallback), and the host-only tail ml_info_to_sqrt is correct."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, _has_cuda


def header_symbols():
    src = open(os.path.join(ROOT, "include", "mins_lidar.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ml_[a-z_0-9]+)\s*\(", src)))


def test_library_loads_and_exports_every_declared_symbol():
    from mins_b200 import capi
    lib = capi.load()
    declared = header_symbols()
    assert len(declared) >= 28
    assert sorted(capi.SYMBOLS) == declared, "capi.SYMBOLS and include/mins_lidar.h disagree"
    for s in declared:
        assert hasattr(lib, s), f"libmins_lidar.so does not export {s}"
    # and nothing torch- or C++-typed leaks through the boundary: every exported ml_* symbol is unmangled
    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(set(re.findall(r"\bT (ml_[a-z_0-9]+)$", out, flags=re.M)))
    assert exported == declared


def test_struct_layouts_match_header(po):
    from mins_b200 import capi
    lib = capi.load()
    sizes = (C.c_int * 4)()
    lib.ml_abi_sizes(sizes)
    assert list(sizes) == [C.sizeof(capi.ml_scan_consts), C.sizeof(capi.ml_point_row), C.sizeof(capi.ml_counters),
                           C.sizeof(capi.ml_config)]
    L = po.lib()
    assert L.orc_sizeof_consts() == sizes[0] and L.orc_sizeof_row() == sizes[1] and L.orc_sizeof_counters() == sizes[2]


def test_header_compiles_as_plain_c(tmp_path):
    """the boundary is a C ABI: the header must be consumable by a C compiler (cgo/JNI/ctypes style bindings)"""
    src = tmp_path / "t.c"
    src.write_text('#include "mins_lidar.h"\nint main(void){ ml_config c; ml_scan_consts s; (void)c; (void)s; return ML_OK; }\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src), "-o",
                           str(tmp_path / "t.o")])


@pytest.mark.skipif(_has_cuda(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback_without_gpu():
    from mins_b200 import capi
    from mins_b200.lidar import LidarMap, MLError
    lib = capi.load()
    cfg = capi.ml_config(0, 1024, 4096, 0, 0.0)
    h = C.c_void_p()
    rc = lib.ml_create(C.byref(cfg), C.byref(h))
    assert rc == capi.ML_ERR_CUDA and not h.value
    assert len(lib.ml_global_error()) > 0
    with pytest.raises(MLError):
        LidarMap(max_scan_points=1024, max_map_points=4096)


def test_missing_library_fails_loudly(monkeypatch):
    from mins_b200 import capi
    monkeypatch.setattr(capi, "_lib", None)
    monkeypatch.setattr(capi, "LIB_PATH", "/nonexistent/libmins_lidar.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        capi.load()


def test_bad_arguments_are_rejected_without_touching_a_device():
    from mins_b200 import capi
    lib = capi.load()
    assert lib.ml_create(None, None) == capi.ML_ERR_ARG
    cfg = capi.ml_config(0, 0, 0, 0, 0.0)
    h = C.c_void_p()
    assert lib.ml_create(C.byref(cfg), C.byref(h)) == capi.ML_ERR_ARG
    assert lib.ml_map_reset(None) == capi.ML_ERR_ARG
    assert lib.ml_map_initialized(None) == 0
    assert lib.ml_launch_count(None) == 0
    lib.ml_destroy(None)


@pytest.mark.parametrize("n,rank", [(6, 6), (48, 6), (48, 48), (55, 20), (1, 1)])
def test_info_to_sqrt_matches_the_information_form(n, rank):
    """Rc^T Rc = HtH and Rc^T rc = Htr for full-rank and rank-deficient systems (planar scenes give rank 6)."""
    from mins_b200.lidar import info_to_sqrt
    rng = np.random.default_rng(n * 100 + rank)
    H = rng.normal(size=(3 * max(rank, 1) + 5, rank)) @ rng.normal(size=(rank, n))
    r = rng.normal(size=len(H))
    HtH, Htr = H.T @ H, H.T @ r
    Rc, rc = info_to_sqrt(HtH, Htr)
    assert len(rc) == rank
    assert np.abs(Rc.T @ Rc - HtH).max() <= 1e-10 * np.abs(HtH).max()
    assert np.abs(Rc.T @ rc - Htr).max() <= 1e-9 * max(np.abs(Htr).max(), 1e-300)


def test_info_to_sqrt_gives_the_same_ekf_update_as_givens_compression(po):
    """StateHelper::measurement_compress_inplace + EKFUpdate (StateHelper.cpp:631-643,121-200) vs our host tail."""
    from mins_b200.lidar import info_to_sqrt
    rng = np.random.default_rng(5)
    n = 48
    H = rng.normal(size=(600, 6)) @ rng.normal(size=(6, n)) + 1e-3 * rng.normal(size=(600, n))
    r = rng.normal(size=600)
    L = rng.normal(size=(n, n))
    P = 1e-4 * (L @ L.T / n + np.eye(n))
    dx0, P0 = po.ekf_update(P, H, r)
    Rc, rc = info_to_sqrt(H.T @ H, H.T @ r)
    dx1, P1 = po.ekf_update(P, Rc, rc)
    assert np.abs(dx1 - dx0).max() <= 1e-9 * np.abs(dx0).max()
    assert np.abs(P1 - P0).max() <= 1e-9 * np.abs(P0).max()
