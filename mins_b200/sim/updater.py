"""UpdaterLidar for the closed-loop harness: the same host logic as mins_b200/host/updater_lidar_gpu.hpp (itself a mirror
of mins/src/update/lidar/UpdaterLidar.cpp:50-408), in Python over a pluggable backend so that the GPU path and the CPU
oracle run in the SAME estimator:

  backend.initialized() / build_map_from_scan(xyz, t_rel, op) / scan_update(xyz, t_rel, op, T, consts, P) -> (rc, Hc, rc_vec, cnt)
  backend.register_scan(xyz, t_rel, op, T) / reanchor(T, now_rel, op)
  backend.last_T : the transform transform_to_map ended up with (the ICP result when op.map_use_icp and ICP succeeded)

GpuBackend (below) drives the C ABI; the oracle backend lives in oracle/backend.py (test infrastructure).
"""
from collections import deque

import numpy as np

from ..lidar import LidarMap, info_to_sqrt, make_scan_consts


class GpuBackend:
    """one LiDAR id = one ml_handle (mins_b200.lidar.LidarMap)."""

    def __init__(self, max_scan_points=1 << 16, max_map_points=1 << 20, device=0):
        self.m = LidarMap(max_scan_points=max_scan_points, max_map_points=max_map_points, device=device)
        self.device = device
        self.resident = None
        self.device_ms = 0.0
        self.last_T = None
        self.last_icp_success = True
        self.icp_runs = self.icp_ok = self.icp_iters = 0

    def close(self):
        self.m.close()

    def _icp(self, scan, op, T_init):
        """the ICP block of transform_to_map (LidarHelper.cpp:156-187) on the freshly uploaded scan cloud"""
        self.m.feed_scan(scan["xyz"], scan["t_rel"], op)
        if op.raw_do_downsample:
            self.m.downsample(op.raw_downsample_size)
        T, res = self.m.icp(T_init, op.map_icp_dist, seed=scan["key"] + 1000003 * scan.get("icp_tries", 0))
        scan["icp_tries"] = scan.get("icp_tries", 0) + 1
        self.icp_runs += 1
        self.icp_ok += int(res.success)
        self.icp_iters += res.iterations
        self.resident = None  # uploaded, not transformed
        return T, bool(res.success)

    def align(self, scan, op, T_init):
        """register_scan's retry of transform_to_map (LidarHelper.cpp:198): returns (T_LtoM, icp_success)"""
        if not op.map_use_icp:
            return np.array(T_init, np.float32).reshape(4, 4), True
        return self._icp(scan, op, T_init)

    def initialized(self):
        return self.m.initialized()

    def build_map_from_scan(self, scan, op, downsample=True):
        self.m.feed_scan(scan["xyz"], scan["t_rel"], op)
        if op.raw_do_downsample and downsample:
            self.m.downsample(op.raw_downsample_size)
        self.m.init_map_local()
        self.resident = scan["key"]

    def scan_update(self, scan, op, T_LtoM, consts, P):
        if op.map_use_icp:
            # transform_to_map with the ICP block (LidarHelper.cpp:156-187), then the point loop on the resident scan
            T_LtoM, self.last_icp_success = self._icp(scan, op, T_LtoM)
            self.m.transform_to_map(T_LtoM)
            rc, HtH, Htr, cnt = self.m.linsys(consts, P)
        else:
            self.last_icp_success = True
            rc, HtH, Htr, cnt = self.m.scan_update(scan["xyz"], scan["t_rel"], op, T_LtoM, consts, P)
        self.last_T = np.array(T_LtoM, np.float32).reshape(4, 4)
        self.device_ms += self.m.timing()["total"]
        self.resident = scan["key"]
        if rc != 0:
            return rc, None, None, cnt
        Hc, rcv = info_to_sqrt(HtH, Htr)  # host tail: stand-in for measurement_compress_inplace
        return rc, Hc, rcv, cnt

    def register_scan(self, scan, op, T_LtoM):
        if self.resident != scan["key"]:
            self.m.feed_scan(scan["xyz"], scan["t_rel"], op)
            self.m.transform_to_map(T_LtoM)
            self.resident = scan["key"]
        return self.m.register_scan(op, wait=False)  # like the C++ shim: queued behind the scan's kernels, not waited for

    def reanchor(self, T_old_to_new, now_minus_FT, op):
        return self.m.propagate_map_to_newest_clone(T_old_to_new, now_minus_FT, op, wait=False)

    def map_size(self):
        return self.m.size()


class LidarUpdater:
    """UpdaterLidar (UpdaterLidar.h:44-90) for one LiDAR id."""

    def __init__(self, state, op, backend, lidar_dt=0.0):
        self.state, self.op, self.backend, self.dt = state, op, backend, lidar_dt
        self.FT = -1.0
        self.ikd_time, self.last_up_time = -1.0, -1.0
        self.stack_raw, self.stack_new, self.stack_used = [], [], []
        self.t_hist = deque(maxlen=101)
        self.n_updates = 0
        self.n_skipped_registrations = 0
        self.counters = []
        self._key = 0

    # ---- Initializer::init_lidar_sim (mins/src/init/Initializer.cpp:308-340): the map is the dense ground-truth scan ---------
    def init_map_gt(self, stamp, xyz):
        if self.FT < 0:
            self.FT = stamp
        self._key += 1
        scan = dict(time=stamp, xyz=np.ascontiguousarray(xyz, np.float32), t_rel=float(np.float32(stamp - self.FT)), key=self._key)
        self.backend.build_map_from_scan(scan, self.op, downsample=False)  # LiDARData ctor + init_map_local, no down-sampling
        self.ikd_time = self.last_up_time = stamp

    # ---- UpdaterLidar::feed_measurement (UpdaterLidar.cpp:50-104) ----------------------------------------------------
    def feed_measurement(self, stamp, xyz):
        st = self.state
        if self.FT < 0:
            self.FT = stamp
        self.t_hist.append(stamp)
        self._key += 1
        self.stack_raw.append(dict(time=stamp, xyz=np.ascontiguousarray(xyz, np.float32), t_rel=float(np.float32(stamp - self.FT)),
                                   key=self._key))
        keep = []
        for scan in self.stack_raw:
            if not st.clones:
                keep.append(scan)
                continue
            t = scan["time"] + self.dt
            if t < st.oldest_clone_time():
                continue
            if t + 0.05 > st.newest_clone_time():
                keep.append(scan)
                continue
            if not self.backend.initialized() or self.last_up_time + 1 < scan["time"]:
                self.backend.build_map_from_scan(scan, self.op)  # LidarHelper::init_map_local
                self.ikd_time = self.last_up_time = scan["time"]
                continue
            self.stack_new.append(scan)
        self.stack_raw = keep
        self._erase_invalid()

    def _erase_invalid(self):
        lim = self.state.oldest_clone_time()
        for name in ("stack_raw", "stack_new", "stack_used"):
            setattr(self, name, [s for s in getattr(self, name) if not (s["time"] + self.dt < lim)])

    # ---- UpdaterLidar::try_update (UpdaterLidar.cpp:106-125) ---------------------------------------------------------
    def try_update(self):
        still = []
        for scan in self.stack_new:
            if self.update(scan):
                self.stack_used.append(scan)
            else:
                still.append(scan)
        self.stack_new = still
        for scan in self.stack_used:
            # LidarHelper::register_scan (LidarHelper.cpp:196-207): "If ICP failed, retry it here" with the state the update has
            # just improved; the scan enters the map - and resets the 1 s staleness clock - only if the alignment succeeded
            if not scan.get("icp_success", True):
                T0 = self._T_LtoM_from_state(scan)
                if T0 is not None:
                    scan["T_LtoM"], scan["icp_success"] = self.backend.align(scan, self.op, T0)
            if scan.get("icp_success", True):
                self.backend.register_scan(scan, self.op, scan["T_LtoM"])
                self.last_up_time = scan["time"]
            else:
                self.n_skipped_registrations += 1
        self.stack_used = []

    def _T_LtoM_from_state(self, scan):
        """the pose algebra of LidarHelper::transform_to_map (LidarHelper.cpp:115-151) from the current estimate"""
        st = self.state
        eI, eA = st.get_interpolated_pose(scan["time"] + self.dt), st.get_interpolated_pose(self.ikd_time + self.dt)
        if eI is None or eA is None:
            return None
        RItoL, pLinI = st.R_ItoL, -st.R_ItoL.T @ st.p_IinL
        RGtoL, pLinG = RItoL @ eI[0], eI[1] + eI[0].T @ pLinI
        RGtoM, pMinG = RItoL @ eA[0], eA[1] + eA[0].T @ pLinI
        T = np.eye(4, dtype=np.float32)
        T[:3, :3] = (RGtoM @ RGtoL.T).astype(np.float32)
        T[:3, 3] = (RGtoM @ (pLinG - pMinG)).astype(np.float32)
        return T

    # ---- UpdaterLidar::propagate_map_frame (UpdaterLidar.cpp:147-160) + propagate_map_to_newest_clone ---------------
    def propagate_map_frame(self):
        st = self.state
        if not self.backend.initialized() or len(st.clones) < st.intr_order + 1:
            return False
        if not (self.ikd_time + self.dt <= st.oldest_3rd_clone_time()):
            return False
        old = st.get_interpolated_pose(self.ikd_time + self.dt)
        if old is None:
            raise RuntimeError("propagate_map_to_newest_clone: cannot get the bounding poses for the map")  # LidarHelper.cpp:244
        tn = st.newest_clone_time()
        new = (st.clones[tn].R, st.clones[tn].p)
        RItoL, pLinI = st.R_ItoL, -st.R_ItoL.T @ st.p_IinL
        RGtoLo, pLo = RItoL @ old[0], old[1] + old[0].T @ pLinI
        RGtoLn, pLn = RItoL @ new[0], new[1] + new[0].T @ pLinI
        T = np.eye(4)
        T[:3, :3] = RGtoLn @ RGtoLo.T
        T[:3, 3] = RGtoLn @ (pLo - pLn)
        self.backend.reanchor(T, st.time - self.FT, self.op)
        self.ikd_time = tn - self.dt
        return True

    # ---- UpdaterLidar::update (UpdaterLidar.cpp:163-408) --------------------------------------------------------------
    def update(self, scan):
        st, op = self.state, self.op
        H_order, map_hx, total_hx = [], {}, 0

        def insert_map(var):
            nonlocal total_hx
            if var not in map_hx:
                map_hx[var] = total_hx
                H_order.append(var)
                total_hx += 6

        if op.do_calib_ext:
            insert_map(("calib",))
        tI, tA = scan["time"] + self.dt, self.ikd_time + self.dt
        jI = st.get_interpolated_jacobian(tI)
        if jI is None:
            return False
        for t, _ in jI[2]:
            insert_map(("clone", t))
        jA = st.get_interpolated_jacobian(tA)
        if jA is None:
            if tA > st.oldest_clone_time():
                return False
            raise RuntimeError("update: cannot get the bounding poses for the map")  # UpdaterLidar.cpp:210
        for t, _ in jA[2]:
            insert_map(("clone", t))
        eI, eA = st.get_interpolated_pose(tI), st.get_interpolated_pose(tA)
        dTdxI = [(map_hx[("clone", t)], D) for t, D in jI[2]]
        dTdxA = [(map_hx[("clone", t)], D) for t, D in jA[2]]
        consts, T_LtoM = make_scan_consts(op, eI[0], eI[1], eA[0], eA[1], jI[0], jI[1], jA[0], jA[1], st.R_ItoL, st.p_IinL, dTdxI,
                                          dTdxA, map_hx[("calib",)] if op.do_calib_ext else -1, total_hx,
                                          at_clone_m=st.have_clone(tI), at_clone_a=st.have_clone(tA), ie_o=st.intr_err_ori,
                                          ie_p=st.intr_err_pos)
        P = st.get_marginal_covariance(H_order)
        rc, Hc, rcv, cnt = self.backend.scan_update(scan, op, T_LtoM, consts, P)
        self.counters.append(cnt)
        if rc != 0 or Hc is None or len(rcv) < 1:
            return False
        if not st.ekf_update(H_order, Hc, rcv):
            return False
        scan["T_LtoM"] = getattr(self.backend, "last_T", None) if op.map_use_icp else T_LtoM  # lidar->T_LtoM
        scan["icp_success"] = bool(getattr(self.backend, "last_icp_success", True))              # lidar->icp_success
        if scan["T_LtoM"] is None:
            scan["T_LtoM"] = T_LtoM
        self.n_updates += 1
        return True
