"""Host-side mirror (Python) of the reference's LiDAR-update interface over the C ABI.

Names follow mins/src/update/lidar/{UpdaterLidar,LidarHelper,LidarTypes}.h and mins/src/options/OptionsLidar.h so
the parity tests read like the reference. The compiled-language mirror (what a MINS maintainer links) is
mins_b200/host/*.h; this module is what tests/ and bench.py drive. All numerics happen in libmins_lidar.so.
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import capi

CHI2_95_3DOF = 7.814727903251179  # boost::math::quantile(chi_squared(3), 0.95), UpdaterStatistics.cpp:31-33


@dataclass
class OptionsLidar:
    """mins/src/options/OptionsLidar.h:38-99 (the knobs the hot path reads); defaults = config/simulation."""
    max_range: float = 100.0
    min_range: float = 1.0
    chi2_mult: float = 1.0
    raw_do_downsample: bool = True
    raw_downsample_size: float = 0.3
    raw_noise: float = 0.01
    raw_remove_motion_blur: bool = False
    raw_point_dt: float = 1e-6
    map_do_downsample: bool = True
    map_downsample_size: float = 0.3
    map_noise: float = 0.5
    map_ngbr_num: int = 5
    map_ngbr_max_d: float = 1.0
    map_decay_time: float = 120.0
    map_decay_dist: float = 100.0
    map_use_icp: bool = False    # reference default true; its ICP is non-deterministic
    map_icp_dist: float = 50.0   # config/simulation/config_lidar.yaml:26
    plane_max_p2pd: float = 0.1
    plane_max_condi: float = 200.0
    do_calib_ext: bool = False
    do_calib_dt: bool = False


def kaist_options(**kw):
    """mins/config/kaist/config_lidar.yaml:11-29 (KAIST Urban: 2 x VLP-16, coarse raw leaf, 0.5 m map, wide neighbour gate)."""
    d = dict(max_range=30.0, min_range=1.0, chi2_mult=1.0, raw_do_downsample=True, raw_downsample_size=2.0, raw_noise=0.1,
             map_do_downsample=True, map_downsample_size=0.5, map_noise=0.1, map_ngbr_num=5, map_ngbr_max_d=10.0, map_decay_time=30.0,
             map_decay_dist=30.0, map_use_icp=False, map_icp_dist=30.0, plane_max_p2pd=0.1, plane_max_condi=200.0)
    d.update(kw)
    return OptionsLidar(**d)


class MLError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"mins_lidar status {code}: {msg}")
        self.code = code


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def make_scan_consts(op, RGtoI_est, pIinG_est, RGtoA_est, pAinG_est, RGtoI_fej, pIinG_fej, RGtoA_fej, pAinG_fej, RItoL,
                     pIinL, dTdxI, dTdxA, calib_col, total_hx, at_clone_m=False, at_clone_a=False, ie_o=0.0, ie_p=0.0):
    """The "Precomputed Transformations" block of UpdaterLidar::update (UpdaterLidar.cpp:223-254).

    dTdxI / dTdxA: lists of (col, D) with D a 6 x width array (State.cpp:745-816). Returns (ml_scan_consts, T_LtoM fp32
    4x4 as LidarHelper::transform_to_map builds it, LidarHelper.cpp:141-151).
    """
    RItoL = _f64(RItoL)
    RLtoI = RItoL.T
    pIinL = _f64(pIinL)
    pLinI = -RLtoI @ pIinL
    c = capi.ml_scan_consts()

    def put(name, a):
        a = _f64(a).ravel()
        getattr(c, name)[:] = a.tolist()

    # est (:233-235, :241-242, :246-247)
    RGtoL_est = RItoL @ RGtoI_est
    pLinG_est = pIinG_est + RGtoI_est.T @ pLinI
    RGtoM_est = RItoL @ RGtoA_est
    pMinG_est = pAinG_est + RGtoA_est.T @ pLinI
    RLtoM_est = RGtoM_est @ RGtoL_est.T
    pLinM_est = RGtoM_est @ (pLinG_est - pMinG_est)
    # fej (:236-239, :243-244, :248-249)
    RItoG_fej = RGtoI_fej.T
    RGtoL_fej = RItoL @ RGtoI_fej
    RLtoG_fej = RGtoL_fej.T
    pLinG_fej = pIinG_fej + RGtoI_fej.T @ pLinI
    RGtoM_fej = RItoL @ RGtoA_fej
    pMinG_fej = pAinG_fej + RGtoA_fej.T @ pLinI
    RLtoM_fej = RGtoM_fej @ RLtoG_fej
    pLinM_fej = RGtoM_fej @ (pLinG_fej - pMinG_fej)
    put("RLtoM_est", RLtoM_est), put("pLinM_est", pLinM_est)
    put("RLtoM_fej", RLtoM_fej), put("pLinM_fej", pLinM_fej)
    put("RGtoM_fej", RGtoM_fej), put("RItoG_fej", RItoG_fej)
    put("RLtoI_fej", RLtoI), put("RItoL_fej", RItoL)
    put("RGtoA_fej", RGtoA_fej), put("RLtoG_fej", RLtoG_fej)
    put("pIinL_fej", pIinL), put("pLinI_fej", pLinI)
    put("pIinG_fej", pIinG_fej), put("pAinG_fej", pAinG_fej)
    c.nI, c.nA = len(dTdxI), len(dTdxA)
    for dst, src in ((c.dTdxI, dTdxI), (c.dTdxA, dTdxA)):
        for j, (col, D) in enumerate(src):
            D = _f64(D).reshape(6, -1)
            dst[j].col, dst[j].width = int(col), D.shape[1]
            buf = np.zeros((6, 6))
            buf[:, :D.shape[1]] = D
            dst[j].D[:] = buf.ravel().tolist()
    c.calib_col = int(calib_col)
    c.at_clone_m, c.at_clone_a = int(at_clone_m), int(at_clone_a)
    c.ie_o, c.ie_p = float(ie_o), float(ie_p)
    c.total_hx = int(total_hx)
    c.raw_noise, c.map_noise = op.raw_noise, op.map_noise
    c.map_ngbr_num = op.map_ngbr_num
    c.map_ngbr_max_d = op.map_ngbr_max_d
    c.plane_max_p2pd, c.plane_max_condi = op.plane_max_p2pd, op.plane_max_condi
    c.chi2_mult = op.chi2_mult  # the library applies quantile(chi_squared(k - 2), 0.95) itself (UpdaterStatistics.cpp:110-111)
    T = np.eye(4, dtype=np.float32)  # LidarHelper.cpp:149-151
    T[:3, :3] = RLtoM_est.astype(np.float32)
    T[:3, 3] = pLinM_est.astype(np.float32)
    return c, T


class LidarMap:
    """One LiDAR id: iKDDATA (LidarTypes.h:71-81) + the LiDARData being processed, backed by one ml_handle."""

    def __init__(self, max_scan_points=262144, max_map_points=1 << 21, grid_cell=0.0, device=0, max_grid_cells=0):
        self.lib = capi.load()
        cfg = capi.ml_config(device, max_scan_points, max_map_points, max_grid_cells, grid_cell)
        h = C.c_void_p()
        rc = self.lib.ml_create(C.byref(cfg), C.byref(h))
        if rc != 0:
            raise MLError(rc, self.lib.ml_global_error().decode())
        self.h = h
        self.time = -1.0          # iKDDATA::time
        self.last_up_time = -1.0  # iKDDATA::last_up_time
        self.hx = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.ml_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, soft_ok=True):
        if rc < 0 or (rc > 0 and not soft_ok):
            raise MLError(rc, self.lib.ml_last_error(self.h).decode())
        return rc

    # ---- KD_TREE surface used through iKDDATA::tree -------------------------------------------------
    def initialized(self):
        return bool(self.lib.ml_map_initialized(self.h))

    def reset(self):
        self._ck(self.lib.ml_map_reset(self.h))

    def Build(self, xyzi):
        a = _f32(xyzi).reshape(-1, 4)
        self._ck(self.lib.ml_map_build(self.h, a.ctypes.data, len(a)))

    def Add_Points(self, xyzi, downsample_on, downsample_size):
        a = _f32(xyzi).reshape(-1, 4)
        n = C.c_int()
        self._ck(self.lib.ml_map_add(self.h, a.ctypes.data, len(a), int(downsample_on), downsample_size, C.byref(n)))
        return n.value

    def flatten(self):
        n = C.c_int()
        self._ck(self.lib.ml_map_size(self.h, C.byref(n)))
        out = np.zeros((n.value, 4), np.float32)
        self._ck(self.lib.ml_map_flatten(self.h, out.ctypes.data, n.value, C.byref(n)))
        return out

    def size(self):
        n = C.c_int()
        self._ck(self.lib.ml_map_size(self.h, C.byref(n)))
        return n.value

    # ---- LiDARData / LidarHelper ---------------------------------------------------------------------
    def feed_scan(self, xyz, t_rel, op):
        """LiDARData ctor (LidarTypes.cpp:26-56). xyz: (n,3) or (n,4) float32. Returns the number kept."""
        a = _f32(xyz)
        n = C.c_int()
        self._ck(self.lib.ml_scan_upload(self.h, a.ctypes.data, a.shape[1], len(a), t_rel, op.min_range, op.max_range, C.byref(n)))
        return n.value

    def set_scan(self, xyzi):
        a = _f32(xyzi).reshape(-1, 4)
        self._ck(self.lib.ml_scan_set(self.h, a.ctypes.data, len(a)))

    def downsample(self, leaf):
        """LidarHelper::downsample (LidarHelper.cpp:89-94)."""
        n = C.c_int()
        self._ck(self.lib.ml_scan_downsample(self.h, leaf, C.byref(n)))
        return n.value

    def deskew(self, knots, total_points):
        a = _f32(knots).reshape(-1, 12)
        self._ck(self.lib.ml_scan_deskew(self.h, a.ctypes.data, len(a), total_points))

    def init_map_local(self):
        """LidarHelper::init_map_local (LidarHelper.cpp:103-113) on the current scan cloud."""
        self._ck(self.lib.ml_map_build_from_scan(self.h))

    def transform_to_map(self, T_LtoM):
        """Tail of LidarHelper::transform_to_map (LidarHelper.cpp:149-151,192)."""
        T = _f32(T_LtoM).reshape(4, 4)
        self._ck(self.lib.ml_scan_transform(self.h, T.ctypes.data))

    def register_scan(self, op, wait=True):
        """LidarHelper::register_scan (LidarHelper.cpp:196-207). wait=False (the reference returns nothing): the merge and the
        grid rebuild are queued and the call returns; the next call on the map picks up the count (and a capacity error)."""
        n = C.c_int()
        self._ck(self.lib.ml_map_add_scan(self.h, int(op.map_do_downsample), op.map_downsample_size, C.byref(n) if wait else None))
        return n.value if wait else None

    def propagate_map_to_newest_clone(self, T_old_to_new, now_minus_FT, op, wait=True):
        """LidarHelper::propagate_map_to_newest_clone (LidarHelper.cpp:209-266), pose look-ups done by the caller."""
        T = _f64(T_old_to_new).reshape(4, 4)
        n = C.c_int()
        self._ck(self.lib.ml_map_reanchor(self.h, T.ctypes.data, now_minus_FT, op.map_decay_time, op.map_decay_dist,
                                          int(op.map_do_downsample), op.map_downsample_size, C.byref(n) if wait else None))
        return n.value if wait else None

    def scan(self, which=0):
        n = C.c_int()
        self._ck(self.lib.ml_scan_download(self.h, which, None, 0, C.byref(n)))
        out = np.zeros((n.value, 4), np.float32)
        if n.value:
            self._ck(self.lib.ml_scan_download(self.h, which, out.ctypes.data, n.value, C.byref(n)))
        return out

    def get_neighbors(self, k, max_d2):
        """LidarHelper::get_neighbors for every point of pointcloud_in_map (LidarHelper.cpp:268-289)."""
        n = C.c_int()
        self._ck(self.lib.ml_scan_download(self.h, 1, None, 0, C.byref(n)))
        idx = np.full((n.value, k), -1, np.int32)
        d2 = np.zeros((n.value, k), np.float32)
        ok = np.zeros(n.value, np.uint8)
        self._ck(self.lib.ml_knn(self.h, k, max_d2, idx.ctypes.data, d2.ctypes.data, ok.ctypes.data, C.byref(n)))
        return idx, d2, ok.astype(bool)

    def linsys(self, consts, P):
        """Point loop + stacking of UpdaterLidar::update (UpdaterLidar.cpp:266-394) -> (status, HtH, Htr, counters)."""
        hx = consts.total_hx
        P = _f64(P).reshape(hx, hx)
        HtH = np.zeros((hx, hx))
        Htr = np.zeros(hx)
        cnt = capi.ml_counters()
        rc = self._ck(self.lib.ml_linsys(self.h, C.byref(consts), P.ctypes.data, HtH.ctypes.data, Htr.ctypes.data, C.byref(cnt)))
        self.hx = hx
        return rc, HtH, Htr, cnt

    def scan_update(self, xyz, t_rel, op, T_LtoM, consts, P, out=None):
        """The timed path in one call: LiDARData ctor -> [downsample] -> transform_to_map -> update()'s point loop."""
        a = xyz if (isinstance(xyz, np.ndarray) and xyz.dtype == np.float32 and xyz.flags.c_contiguous) else _f32(xyz)
        hx = consts.total_hx
        T = _f32(T_LtoM).reshape(4, 4)
        P = _f64(P).reshape(hx, hx)
        if out is None:
            out = (np.zeros((hx, hx)), np.zeros(hx), capi.ml_counters())
        HtH, Htr, cnt = out
        leaf = op.raw_downsample_size if op.raw_do_downsample else 0.0
        rc = self._ck(self.lib.ml_scan_update(self.h, a.ctypes.data, a.shape[1], len(a), t_rel, op.min_range, op.max_range, leaf,
                                              T.ctypes.data, C.byref(consts), P.ctypes.data, HtH.ctypes.data, Htr.ctypes.data,
                                              C.byref(cnt)))
        self.hx = hx
        return rc, HtH, Htr, cnt

    def point_rows(self, want_h=True):
        n = C.c_int()
        self._ck(self.lib.ml_point_rows(self.h, None, None, 0, C.byref(n)))
        rows = (capi.ml_point_row * max(n.value, 1))()
        h = np.zeros((n.value, self.hx)) if want_h else None
        self._ck(self.lib.ml_point_rows(self.h, rows, h.ctypes.data if want_h else None, n.value, C.byref(n)))
        return rows_to_numpy(rows, n.value), h

    def icp(self, T_init, map_icp_dist, seed=0, **overrides):
        """The ICP block of LidarHelper::transform_to_map (LidarHelper.cpp:156-187) on the current scan cloud.
        Returns (T_LtoM fp32 4x4 to pass to transform_to_map, ml_icp_result)."""
        o = capi.ml_icp_options()
        self.lib.ml_icp_default_options(C.byref(o))
        o.max_dist, o.seed = float(map_icp_dist), int(seed)
        for k, v in overrides.items():
            setattr(o, k, v)
        T = _f32(T_init).reshape(4, 4)
        res = capi.ml_icp_result()
        self._ck(self.lib.ml_scan_icp(self.h, T.ctypes.data, C.byref(o), C.byref(res)))
        return np.array(res.T_LtoM, np.float32).reshape(4, 4), res

    # ---- simulator LiDAR model (Simulator::get_lidar_pointcloud, Simulator.cpp:742-815) --------------
    def sim_set_world(self, planes22):
        a = _f64(planes22).reshape(-1, 22)
        self._ck(self.lib.ml_sim_set_world(self.h, a.ctypes.data, len(a)))

    def sim_set_rays(self, dirs_L):
        a = _f64(dirs_L).reshape(-1, 3)
        self._ck(self.lib.ml_sim_set_rays(self.h, a.ctypes.data, len(a)))
        self._sim_rays = len(a)

    def sim_raycast(self, p_LinG, R_LtoG, min_range, max_range, noise=0.0, seed=0, download=True):
        """Returns the (n, 4) float32 scan (or just n when download=False; the scan then stays on the device only)."""
        p, R = _f64(p_LinG).reshape(3), _f64(R_LtoG).reshape(3, 3)
        n = C.c_int()
        out = np.zeros((self._sim_rays, 4), np.float32) if download else None
        self._ck(self.lib.ml_sim_raycast(self.h, p.ctypes.data, R.ctypes.data, min_range, max_range, noise, int(seed),
                                         out.ctypes.data if download else None, self._sim_rays if download else 0, C.byref(n)))
        return out[:n.value] if download else n.value

    def sim_raycast_sweep(self, pose_knots, min_range, max_range, noise=0.0, seed=0):
        """Motion-distorted scan (Simulator.cpp:761-771). pose_knots: (n_knots, 12) [R_LtoG row-major | p_LinG]."""
        k = _f64(pose_knots).reshape(-1, 12)
        n = C.c_int()
        out = np.zeros((self._sim_rays, 4), np.float32)
        self._ck(self.lib.ml_sim_raycast_sweep(self.h, k.ctypes.data, len(k), min_range, max_range, noise, int(seed), out.ctypes.data,
                                               self._sim_rays, C.byref(n)))
        return out[:n.value]

    def sim_scan_device(self):
        """(device pointer, n) of the last ray-cast scan, to pass as `xyz` of ml_scan_update (stride 4)."""
        ptr, n = C.c_void_p(), C.c_int()
        self._ck(self.lib.ml_sim_scan_device(self.h, C.byref(ptr), C.byref(n)))
        return ptr.value, n.value

    def timing(self):
        ms = (C.c_float * 8)()
        self.lib.ml_last_timing(self.h, ms)
        return dict(zip(["upload_filter", "downsample", "transform", "bin", "knn", "rows_reduce", "d2h", "total"], list(ms)))

    def set_stats(self, on):
        """residual statistics of UpdaterStatistics::Chi2Check (ml_counters.sum_res) on/off"""
        self.lib.ml_set_stats(self.h, int(on))

    def set_timing(self, on):
        self.lib.ml_set_timing(self.h, int(on))

    def launch_count(self):
        return int(self.lib.ml_launch_count(self.h))


def rows_to_numpy(rows, n):
    dt = np.dtype([("status", np.int32), ("pad", np.int32), ("plane", np.float64, 4), ("res", np.float64),
                   ("chi2", np.float64), ("w", np.float64), ("vr", np.float64), ("raw_n", np.float64)])
    assert dt.itemsize == C.sizeof(capi.ml_point_row)
    return np.frombuffer(rows, dtype=dt, count=n).copy()


def info_to_sqrt(HtH, Htr):
    """Host tail: (HtH, Htr) -> compressed (H_c, r_c) for StateHelper::EKFUpdate (StateHelper.cpp:631-643 stand-in)."""
    lib = capi.load()
    n = len(Htr)
    HtH, Htr = _f64(HtH), _f64(Htr)
    Rc = np.zeros((n, n))
    rc = np.zeros(n)
    rank = lib.ml_info_to_sqrt(HtH.ctypes.data, Htr.ctypes.data, n, Rc.ctypes.data, rc.ctypes.data)
    if rank < 0:
        raise MLError(rank, "ml_info_to_sqrt")
    return Rc[:rank], rc[:rank]
