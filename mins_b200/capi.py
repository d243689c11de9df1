"""ctypes binding of include/mins_lidar.h (the C-ABI shared library mins_b200/libmins_lidar.so).

The library is hand-written CUDA for sm_100a; there is NO CPU fallback. ``load()`` raises if the shared library is
missing (run ``python -c 'import __graft_entry__ as g; g.build()'`` first) and ``ml_create`` fails if no CUDA device
is usable.
"""
import ctypes as C
import os

ML_MAX_HX = 64
ML_MAX_BLOCKS = 8
ML_MAX_K = 10
ML_OK, ML_SOFT_NO_ROWS, ML_SOFT_NO_MAP = 0, 1, 2
ML_ERR_CUDA, ML_ERR_CAPACITY, ML_ERR_ARG = -1, -2, -3

_HERE = os.path.dirname(os.path.abspath(__file__))
# MINS_LIDAR_LIB: development override to A/B-test a differently compiled build of the same library
LIB_PATH = os.environ.get("MINS_LIDAR_LIB") or os.path.join(_HERE, "libmins_lidar.so")


class ml_config(C.Structure):
    _fields_ = [("device", C.c_int), ("max_scan_points", C.c_int), ("max_map_points", C.c_int),
                ("max_grid_cells", C.c_int), ("grid_cell", C.c_float)]


class ml_jac_block(C.Structure):
    _fields_ = [("col", C.c_int), ("width", C.c_int), ("D", C.c_double * 36)]


class ml_scan_consts(C.Structure):
    _fields_ = [
        ("RLtoM_est", C.c_double * 9), ("pLinM_est", C.c_double * 3),
        ("RLtoM_fej", C.c_double * 9), ("pLinM_fej", C.c_double * 3),
        ("RGtoM_fej", C.c_double * 9), ("RItoG_fej", C.c_double * 9),
        ("RLtoI_fej", C.c_double * 9), ("RItoL_fej", C.c_double * 9),
        ("RGtoA_fej", C.c_double * 9), ("RLtoG_fej", C.c_double * 9),
        ("pIinL_fej", C.c_double * 3), ("pLinI_fej", C.c_double * 3),
        ("pIinG_fej", C.c_double * 3), ("pAinG_fej", C.c_double * 3),
        ("nI", C.c_int), ("nA", C.c_int),
        ("dTdxI", ml_jac_block * ML_MAX_BLOCKS), ("dTdxA", ml_jac_block * ML_MAX_BLOCKS),
        ("calib_col", C.c_int), ("at_clone_m", C.c_int), ("at_clone_a", C.c_int),
        ("ie_o", C.c_double), ("ie_p", C.c_double), ("total_hx", C.c_int),
        ("raw_noise", C.c_double), ("map_noise", C.c_double), ("map_ngbr_num", C.c_int),
        ("map_ngbr_max_d", C.c_double), ("plane_max_p2pd", C.c_double), ("plane_max_condi", C.c_double),
        ("chi2_mult", C.c_double),
    ]


class ml_counters(C.Structure):
    _fields_ = [("n_points", C.c_int), ("n_knn_fail", C.c_int), ("n_plane_fail", C.c_int), ("n_badmatch", C.c_int),
                ("n_chi_rej", C.c_int), ("n_acc", C.c_int), ("sum_chi", C.c_double), ("sum_chi2", C.c_double),
                ("sum_res", C.c_double), ("sum_res2", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class ml_point_row(C.Structure):
    _fields_ = [("status", C.c_int), ("pad", C.c_int), ("plane", C.c_double * 4), ("res", C.c_double),
                ("chi2", C.c_double), ("w", C.c_double), ("vr", C.c_double), ("raw_n", C.c_double)]


class ml_icp_options(C.Structure):
    _fields_ = [("max_dist", C.c_double), ("read_prob", C.c_double), ("ref_ratio", C.c_double), ("normal_knn", C.c_int),
                ("trim_ratio", C.c_double), ("max_iter", C.c_int), ("min_diff_rot", C.c_double), ("min_diff_trans", C.c_double),
                ("smooth_len", C.c_int), ("max_match_dist", C.c_double), ("normal_radius", C.c_double), ("seed", C.c_uint64)]


class ml_icp_result(C.Structure):
    _fields_ = [("T_LtoM", C.c_float * 16), ("T_icp", C.c_float * 16), ("success", C.c_int), ("iterations", C.c_int),
                ("max_iter_reached", C.c_int), ("n_reading", C.c_int), ("n_reference", C.c_int), ("n_matched", C.c_int),
                ("n_used", C.c_int), ("mean_sq_err", C.c_double)]


# every symbol include/mins_lidar.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "ml_create", "ml_destroy", "ml_last_error", "ml_global_error", "ml_map_reset", "ml_map_initialized", "ml_map_build",
    "ml_map_build_from_scan", "ml_map_add", "ml_map_add_scan", "ml_map_reanchor", "ml_map_flatten", "ml_map_size",
    "ml_scan_upload", "ml_scan_deskew", "ml_scan_downsample", "ml_scan_transform", "ml_scan_download", "ml_scan_set",
    "ml_knn", "ml_linsys", "ml_scan_update", "ml_point_rows", "ml_info_to_sqrt", "ml_last_timing", "ml_set_timing",
    "ml_abi_sizes", "ml_launch_count", "ml_sim_set_world", "ml_sim_set_rays", "ml_sim_raycast", "ml_sim_raycast_sweep", "ml_sim_scan_device",
    "ml_icp_default_options", "ml_scan_icp", "ml_debug_timeline", "ml_host_alloc", "ml_host_free", "ml_set_stats",
    "ml_range_filter_static",
]

_lib = None


def load():
    """dlopen the in-tree CUDA library. Fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the CUDA extension has not been built (run __graft_entry__.build()). "
            "mins_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, ip, fp, dp = C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_float), C.POINTER(C.c_double)
    lib.ml_create.argtypes = [C.POINTER(ml_config), C.POINTER(vp)]
    lib.ml_destroy.argtypes = [vp]
    lib.ml_destroy.restype = None
    lib.ml_last_error.argtypes = [vp]
    lib.ml_last_error.restype = C.c_char_p
    lib.ml_global_error.restype = C.c_char_p
    lib.ml_map_reset.argtypes = [vp]
    lib.ml_map_initialized.argtypes = [vp]
    lib.ml_map_build.argtypes = [vp, vp, C.c_int]
    lib.ml_map_build_from_scan.argtypes = [vp]
    lib.ml_map_add.argtypes = [vp, vp, C.c_int, C.c_int, C.c_float, ip]
    lib.ml_map_add_scan.argtypes = [vp, C.c_int, C.c_float, ip]
    lib.ml_map_reanchor.argtypes = [vp, vp, C.c_double, C.c_double, C.c_double, C.c_int, C.c_float, ip]
    lib.ml_map_flatten.argtypes = [vp, vp, C.c_int, ip]
    lib.ml_map_size.argtypes = [vp, ip]
    lib.ml_scan_upload.argtypes = [vp, vp, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, ip]
    lib.ml_scan_deskew.argtypes = [vp, vp, C.c_int, C.c_int]
    lib.ml_scan_downsample.argtypes = [vp, C.c_float, ip]
    lib.ml_scan_transform.argtypes = [vp, vp]
    lib.ml_scan_download.argtypes = [vp, C.c_int, vp, C.c_int, ip]
    lib.ml_scan_set.argtypes = [vp, vp, C.c_int]
    lib.ml_knn.argtypes = [vp, C.c_int, C.c_float, vp, vp, vp, ip]
    lib.ml_linsys.argtypes = [vp, C.POINTER(ml_scan_consts), vp, vp, vp, C.POINTER(ml_counters)]
    lib.ml_scan_update.argtypes = [vp, vp, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, vp,
                                   C.POINTER(ml_scan_consts), vp, vp, vp, C.POINTER(ml_counters)]
    lib.ml_point_rows.argtypes = [vp, vp, vp, C.c_int, ip]
    lib.ml_info_to_sqrt.argtypes = [vp, vp, C.c_int, vp, vp]
    lib.ml_last_timing.argtypes = [vp, vp]
    lib.ml_set_timing.argtypes = [vp, C.c_int]
    lib.ml_abi_sizes.argtypes = [vp]
    lib.ml_abi_sizes.restype = None
    lib.ml_launch_count.argtypes = [vp]
    lib.ml_launch_count.restype = C.c_longlong
    lib.ml_icp_default_options.argtypes = [C.POINTER(ml_icp_options)]
    lib.ml_icp_default_options.restype = None
    lib.ml_scan_icp.argtypes = [vp, vp, C.POINTER(ml_icp_options), C.POINTER(ml_icp_result)]
    lib.ml_debug_timeline.argtypes = [vp, C.c_int, vp, C.c_int]
    lib.ml_set_stats.argtypes = [vp, C.c_int]
    lib.ml_host_alloc.argtypes = [C.c_size_t]
    lib.ml_host_alloc.restype = vp
    lib.ml_host_free.argtypes = [vp]
    lib.ml_host_free.restype = None
    lib.ml_range_filter_static.argtypes = [vp, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, vp, vp, vp]
    lib.ml_sim_set_world.argtypes = [vp, vp, C.c_int]
    lib.ml_sim_set_rays.argtypes = [vp, vp, C.c_int]
    lib.ml_sim_raycast.argtypes = [vp, vp, vp, C.c_double, C.c_double, C.c_float, C.c_uint64, vp, C.c_int, ip]
    lib.ml_sim_raycast_sweep.argtypes = [vp, vp, C.c_int, C.c_double, C.c_double, C.c_float, C.c_uint64, vp, C.c_int, ip]
    lib.ml_sim_scan_device.argtypes = [vp, C.POINTER(vp), ip]
    sizes = (C.c_int * 4)()
    lib.ml_abi_sizes(sizes)
    want = [C.sizeof(ml_scan_consts), C.sizeof(ml_point_row), C.sizeof(ml_counters), C.sizeof(ml_config)]
    if list(sizes) != want:
        raise RuntimeError(f"ABI layout mismatch between capi.py {want} and libmins_lidar.so {list(sizes)}")
    _lib = lib
    return lib
