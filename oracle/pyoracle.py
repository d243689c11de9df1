"""TEST INFRASTRUCTURE ONLY — ctypes access to the CPU oracle (oracle/liboracle.so: restatement; oracle/_ref/
libikd_ref.so: the reference's own ikd-Tree compiled unmodified). Imported only by tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs. The product package (mins_b200/) never imports this.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from mins_b200 import capi  # struct layouts only (shared public header)  # noqa: E402

_lib = None
_ref = None


def build(force=False):
    """make -C oracle (restatement always; _ref only where /root/reference exists)."""
    if force or not os.path.exists(os.path.join(_HERE, "liboracle.so")) or (
            os.path.exists("/root/reference/mins/src/update/lidar/ikd_Tree.cpp")
            and not os.path.exists(os.path.join(_HERE, "_ref", "libikd_ref.so"))):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(os.path.join(_HERE, "liboracle.so"))
        vp = C.c_void_p
        L.orc_knn_brute.argtypes = [vp, C.c_int, vp, C.c_int, C.c_int, C.c_float, vp, vp, vp, C.c_int]
        L.orc_knn_brute.restype = None
        L.orc_compute_plane.argtypes = [vp, C.c_int, C.c_double, C.c_double, vp, vp]
        L.orc_point.argtypes = [C.POINTER(capi.ml_scan_consts), vp, vp, vp, vp, vp, vp, vp]
        L.orc_scan_linsys.argtypes = [C.POINTER(capi.ml_scan_consts), vp, vp, C.c_int, vp, vp, C.c_int, vp, vp, vp,
                                      C.POINTER(capi.ml_counters), vp, vp, vp, vp, vp, C.c_int]
        L.orc_ekf_update.argtypes = [vp, C.c_int, vp, vp, C.c_int, vp, vp]
        L.orc_range_filter.argtypes = [vp, C.c_int, C.c_int, C.c_float, C.c_double, C.c_double, vp]
        L.orc_transform_f32.argtypes = [vp, C.c_int, vp, vp]
        L.orc_transform_f32.restype = None
        L.orc_transform_f64.argtypes = [vp, C.c_int, vp, vp]
        L.orc_transform_f64.restype = None
        L.orc_voxel_downsample.argtypes = [vp, C.c_int, C.c_float, vp, vp]
        L.orc_map_insert.argtypes = [vp, C.c_int, vp, C.c_int, C.c_int, C.c_float, vp, C.c_int]
        L.orc_reanchor.argtypes = [vp, C.c_int, vp, C.c_double, C.c_double, C.c_double, C.c_int, C.c_float, vp]
        L.orc_deskew.argtypes = [vp, C.c_int, vp]
        L.orc_deskew.restype = None
        L.orc_chi2_quantile95.argtypes = [C.c_int]
        L.orc_chi2_quantile95.restype = C.c_double
        L.orc_scan_linsys_nb.argtypes = [C.POINTER(capi.ml_scan_consts), vp, vp, C.c_int, vp, vp, vp, vp, C.POINTER(capi.ml_counters),
                                         vp, vp, vp, C.c_int]
        assert L.orc_sizeof_consts() == C.sizeof(capi.ml_scan_consts)
        assert L.orc_sizeof_row() == C.sizeof(capi.ml_point_row)
        assert L.orc_sizeof_counters() == C.sizeof(capi.ml_counters)
        _lib = L
    return _lib


def ref_available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libikd_ref.so"))


def ref():
    """the reference's own ikd-Tree (None if the prebuilt .so is absent)."""
    global _ref
    if _ref is None:
        build()
        p = os.path.join(_HERE, "_ref", "libikd_ref.so")
        if not os.path.exists(p):
            return None
        L = C.CDLL(p)
        vp = C.c_void_p
        L.ref_ikd_create.argtypes = [C.c_float]
        L.ref_ikd_create.restype = vp
        L.ref_ikd_destroy.argtypes = [vp]
        L.ref_ikd_destroy.restype = None
        L.ref_ikd_build.argtypes = [vp, vp, C.c_int]
        L.ref_ikd_build.restype = None
        L.ref_ikd_size.argtypes = [vp]
        L.ref_ikd_add.argtypes = [vp, vp, C.c_int, C.c_int]
        L.ref_ikd_flatten.argtypes = [vp, vp, C.c_int]
        L.ref_ikd_knn.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, vp, C.c_int]
        L.ref_ikd_knn.restype = None
        L.ref_lidar_update.argtypes = [vp, vp, C.c_int, C.c_int, C.c_float, C.c_double, C.c_double, vp,
                                       C.POINTER(capi.ml_scan_consts), vp, vp, vp, vp, C.POINTER(capi.ml_counters), C.c_int, vp, vp]
        _ref = L
    return _ref


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def knn_brute(map_xyzi, q_xyzi, k, max_d2, threads=8):
    m, q = f32(map_xyzi).reshape(-1, 4), f32(q_xyzi).reshape(-1, 4)
    idx = np.zeros((len(q), k), np.int32)
    d2 = np.zeros((len(q), k), np.float32)
    ok = np.zeros(len(q), np.uint8)
    lib().orc_knn_brute(m.ctypes.data, len(m), q.ctypes.data, len(q), k, max_d2, idx.ctypes.data, d2.ctypes.data,
                        ok.ctypes.data, threads)
    return idx, d2, ok.astype(bool)


def compute_plane(nb_xyzi, max_condi=200.0, max_p2pd=0.1):
    nb = f32(nb_xyzi).reshape(-1, 4)
    plane = np.zeros(4)
    cond = C.c_double()
    ok = lib().orc_compute_plane(nb.ctypes.data, len(nb), max_condi, max_p2pd, plane.ctypes.data, C.byref(cond))
    return bool(ok), plane, cond.value


def scan_linsys(consts, P, map_xyzi, scan_L, scan_M, knn_idx=None, threads=8, want_compress=False):
    from mins_b200.lidar import rows_to_numpy
    hx = consts.total_hx
    m, sl, sm = f32(map_xyzi).reshape(-1, 4), f32(scan_L).reshape(-1, 4), f32(scan_M).reshape(-1, 4)
    P = f64(P)
    N = len(sl)
    HtH, Htr = np.zeros((hx, hx)), np.zeros(hx)
    cnt = capi.ml_counters()
    rows = (capi.ml_point_row * max(N, 1))()
    hrows = np.zeros((N, hx))
    Hc, rc, nc = np.zeros((hx, hx)), np.zeros(hx), C.c_int(0)
    ki = None if knn_idx is None else np.ascontiguousarray(knn_idx, np.int32)
    total = lib().orc_scan_linsys(C.byref(consts), P.ctypes.data, m.ctypes.data, len(m), sl.ctypes.data, sm.ctypes.data, N,
                                  None if ki is None else ki.ctypes.data, HtH.ctypes.data, Htr.ctypes.data, C.byref(cnt), rows,
                                  hrows.ctypes.data, Hc.ctypes.data if want_compress else None,
                                  rc.ctypes.data if want_compress else None, C.byref(nc) if want_compress else None, threads)
    out = dict(HtH=HtH, Htr=Htr, cnt=cnt, rows=rows_to_numpy(rows, N), h=hrows, total_res=total)
    if want_compress:
        out["Hc"], out["rc"] = Hc[:nc.value], rc[:nc.value]
    return out


def chi2_quantile95(dof):
    """quantile(chi_squared(dof), 0.95) as the oracle computes it (UpdaterStatistics.cpp:31-33)."""
    return float(lib().orc_chi2_quantile95(int(dof)))


def scan_linsys_nb(consts, P, scan_L, nb_xyzi, nb_ok, threads=8, want_compress=False):
    """rows -> stacking [-> Givens compression] driven by neighbour COORDINATES (N x k x 4 float32), the form in which
    KD_TREE::Nearest_Search returns them, with nb_ok = LidarHelper::get_neighbors' verdict per point."""
    hx = consts.total_hx
    sl = f32(scan_L).reshape(-1, 4)
    nb = f32(nb_xyzi).reshape(len(sl), -1, 4)
    ok = np.ascontiguousarray(nb_ok, np.uint8)
    P = f64(P)
    HtH, Htr = np.zeros((hx, hx)), np.zeros(hx)
    cnt = capi.ml_counters()
    Hc, rc, nc = np.zeros((hx, hx)), np.zeros(hx), C.c_int(0)
    total = lib().orc_scan_linsys_nb(C.byref(consts), P.ctypes.data, sl.ctypes.data, len(sl), nb.ctypes.data, ok.ctypes.data,
                                     HtH.ctypes.data, Htr.ctypes.data, C.byref(cnt), Hc.ctypes.data if want_compress else None,
                                     rc.ctypes.data if want_compress else None, C.byref(nc) if want_compress else None, threads)
    out = dict(HtH=HtH, Htr=Htr, cnt=cnt, total_res=total)
    if want_compress:
        out["Hc"], out["rc"] = Hc[:nc.value], rc[:nc.value]
    return out


def ekf_update(P, H, res):
    P, H, res = f64(P), f64(H), f64(res)
    n, m = P.shape[0], H.shape[0]
    dx, Pn = np.zeros(n), np.zeros((n, n))
    rcode = lib().orc_ekf_update(P.ctypes.data, n, H.ctypes.data, res.ctypes.data, m, dx.ctypes.data, Pn.ctypes.data)
    assert rcode == 0
    return dx, Pn


def range_filter(xyz, t_rel, max_range, min_range):
    a = f32(xyz)
    out = np.zeros((len(a), 4), np.float32)
    n = lib().orc_range_filter(a.ctypes.data, a.shape[1], len(a), t_rel, max_range, min_range, out.ctypes.data)
    return out[:n]


def transform_f32(xyzi, T):
    a, T = f32(xyzi).reshape(-1, 4), f32(T).reshape(4, 4)
    out = np.zeros_like(a)
    lib().orc_transform_f32(a.ctypes.data, len(a), T.ctypes.data, out.ctypes.data)
    return out


def voxel_downsample(xyzi, leaf):
    a = f32(xyzi).reshape(-1, 4)
    out = np.zeros_like(a)
    vid = np.zeros(len(a), np.uint32)
    n = lib().orc_voxel_downsample(a.ctypes.data, len(a), leaf, out.ctypes.data, vid.ctypes.data)
    return out[:n], vid[:n]


def map_insert(map_xyzi, add_xyzi, downsample, ds):
    m, a = f32(map_xyzi).reshape(-1, 4), f32(add_xyzi).reshape(-1, 4)
    out = np.zeros((len(m) + len(a), 4), np.float32)
    n = lib().orc_map_insert(m.ctypes.data, len(m), a.ctypes.data, len(a), int(downsample), ds, out.ctypes.data, len(out))
    return out[:n]


def reanchor(map_xyzi, T, now_minus_FT, decay_time, decay_dist, do_ds, ds):
    m, T = f32(map_xyzi).reshape(-1, 4), f64(T).reshape(4, 4)
    out = np.zeros_like(m)
    n = lib().orc_reanchor(m.ctypes.data, len(m), T.ctypes.data, now_minus_FT, decay_time, decay_dist, int(do_ds), ds,
                           out.ctypes.data)
    return out[:n]


def deskew(xyzi, T_per_point):
    a = f32(xyzi).reshape(-1, 4).copy()
    T = f32(T_per_point).reshape(-1, 12)
    lib().orc_deskew(a.ctypes.data, len(a), T.ctypes.data)
    return a


# ---- simulator LiDAR model (mins/src/sim/Simulator.cpp:742-815, SimulationPlane.h:43-95) restated in numpy -------------
def philox4x32(counter0, key_lo, key_hi):
    """Philox-4x32-10 (Salmon et al. 2011, the published algorithm) for counters (counter0[i], 0, 0, 0): (n, 4) uint32."""
    c = np.zeros((len(counter0), 4), np.uint64)
    c[:, 0] = np.asarray(counter0, np.uint64)
    k0, k1 = np.uint64(key_lo), np.uint64(key_hi)
    M0, M1, mask = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c[:, 0], M1 * c[:, 2]
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask, p1 >> np.uint64(32), p1 & mask
        c = np.stack([hi1 ^ c[:, 1] ^ k0, lo1, hi0 ^ c[:, 3] ^ k1, lo0], 1)
        k0 = (k0 + np.uint64(0x9E3779B9)) & mask
        k1 = (k1 + np.uint64(0xBB67AE85)) & mask
    return c.astype(np.uint32)


def sim_noise(ray_idx, seed):
    """(n, 3) float32 N(0,1): Box-Muller on Philox words, u = ((x >> 8) + 1) * 2^-24 (the scheme of csrc/sim.cuh)."""
    w = philox4x32(ray_idx, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    u = ((w >> np.uint32(8)).astype(np.float32) + np.float32(1.0)) * np.float32(5.9604644775390625e-08)
    r0 = np.sqrt(np.float32(-2.0) * np.log(u[:, 0]))
    r1 = np.sqrt(np.float32(-2.0) * np.log(u[:, 2]))
    th0, th1 = np.float32(6.283185307179586) * u[:, 1], np.float32(6.283185307179586) * u[:, 3]
    return np.stack([r0 * np.cos(th0), r0 * np.sin(th0), r1 * np.cos(th1)], 1).astype(np.float32)


def sim_lidar_scan(corners, p_LinG, R_LtoG, dirs_L, min_range, max_range, noise=0.0, seed=0):
    """Simulator::get_lidar_pointcloud for one pose. corners: (P, 4, 3) [top_left, top_right, bottom_left, bottom_right]
    (Simulator.cpp:1053-1090). Literal SimulationPlane::calculate_intersection incl. the normalisations. Returns the (n, 4)
    float32 scan, the kept ray indices and the hit distance of every ray."""
    q = f64(corners).reshape(-1, 4, 3)
    tl, tr, bl, br = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    N = np.cross(tr - tl, bl - tl)                      # SimulationPlane.h:46-49
    D = -(N * tl).sum(1)                                # :53
    unit = lambda v: v / np.linalg.norm(v, axis=-1, keepdims=True)
    V1, V2, V3, V4 = unit(tr - tl), unit(bl - tl), unit(tr - br), unit(bl - br)
    o = f64(p_LinG).reshape(3)
    dl = f64(dirs_L).reshape(-1, 3)
    dg = dl @ f64(R_LtoG).reshape(3, 3).T               # Simulator.cpp:769
    best = np.full(len(dg), np.inf)
    with np.errstate(divide="ignore", invalid="ignore"):
        for s in range(0, len(dg), 4096):
            d = dg[s:s + 4096]
            t = -(N @ o + D)[None, :] / (d @ N.T)       # :76
            pt = o[None, None, :] + t[:, :, None] * d[:, None, :]
            U1, U2 = unit(pt - tl[None]), unit(pt - br[None])
            ok = (t > 0) & ((U1 * V1[None]).sum(2) > 0) & ((U1 * V2[None]).sum(2) > 0) & ((U2 * V3[None]).sum(2) > 0) & \
                 ((U2 * V4[None]).sum(2) > 0)           # :94
            best[s:s + 4096] = np.where(ok, t, np.inf).min(1)
    keep = np.where(~((best > max_range) | (best < min_range)))[0]  # Simulator.cpp:788
    pts = np.zeros((len(keep), 4), np.float32)
    pts[:, :3] = (best[keep, None] * dl[keep]).astype(np.float32)  # :792-795
    if noise > 0:
        pts[:, :3] = pts[:, :3] + np.float32(noise) * sim_noise(keep, seed)  # :798-803 (generator: see ml_sim_raycast)
    return pts, keep, best


class RefIkd:
    """The reference's KD_TREE<PointXYZI> (compiled unmodified)."""

    def __init__(self, downsample_size=0.3):
        self.L = ref()
        if self.L is None:
            raise RuntimeError("oracle/_ref/libikd_ref.so absent")
        self.h = self.L.ref_ikd_create(downsample_size)

    def __del__(self):
        try:
            if self.h:
                self.L.ref_ikd_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def build(self, xyzi):
        a = f32(xyzi).reshape(-1, 4)
        self.L.ref_ikd_build(self.h, a.ctypes.data, len(a))

    def add(self, xyzi, downsample=True):
        a = f32(xyzi).reshape(-1, 4)
        return self.L.ref_ikd_add(self.h, a.ctypes.data, len(a), int(downsample))

    def size(self):
        return self.L.ref_ikd_size(self.h)

    def flatten(self):
        n = self.size()
        out = np.zeros((n + 16, 4), np.float32)
        m = self.L.ref_ikd_flatten(self.h, out.ctypes.data, len(out))
        return out[:m]

    def knn(self, q_xyzi, k, threads=8):
        q = f32(q_xyzi).reshape(-1, 4)
        nb = np.zeros((len(q), k, 4), np.float32)
        d2 = np.zeros((len(q), k), np.float32)
        found = np.zeros(len(q), np.int32)
        self.L.ref_ikd_knn(self.h, q.ctypes.data, len(q), k, nb.ctypes.data, d2.ctypes.data, found.ctypes.data, threads)
        return nb, d2, found

    def lidar_update(self, raw, t_rel, op, T_LtoM, consts, P, threads=8, want_info=True):
        """The reference CPU path for one scan: LiDARData ctor -> transform -> reference ikd-Tree kNN in
        the static-chunk thread loop -> restated rows -> stacking -> Givens compression. Returns (Hc, rc, counters, HtH, Htr).
        want_info=False (the timed baseline): HtH / Htr are NOT formed - the reference never forms them (UpdaterLidar.cpp:376-397
        stacks and QR-compresses); they exist only so that tests can compare with the device's information form."""
        a = f32(raw)
        hx = consts.total_hx
        T, P = f32(T_LtoM).reshape(4, 4), f64(P)
        Hc, rc, nc = np.zeros((hx, hx)), np.zeros(hx), C.c_int(0)
        HtH, Htr = np.zeros((hx, hx)), np.zeros(hx)
        cnt = capi.ml_counters()
        self.L.ref_lidar_update(self.h, a.ctypes.data, a.shape[1], len(a), t_rel, op.max_range, op.min_range, T.ctypes.data,
                                C.byref(consts), P.ctypes.data, Hc.ctypes.data, rc.ctypes.data, C.byref(nc), C.byref(cnt), threads,
                                HtH.ctypes.data if want_info else None, Htr.ctypes.data if want_info else None)
        return Hc[:nc.value], rc[:nc.value], cnt, HtH, Htr
