"""TEST INFRASTRUCTURE ONLY — CPU-oracle backend for the closed-loop harness (mins_b200/sim/updater.py).

Same interface as mins_b200.sim.updater.GpuBackend, every step done by the CPU oracle (oracle/pyoracle.py): range filter,
PCL-VoxelGrid restatement, fp32 transform, brute-force exact kNN, restated rows, stacking and the reference's Givens
measurement compression (StateHelper.cpp:631-643), map insert and re-anchor. Used by tests/ to compare closed-loop
trajectories (GPU updater vs CPU oracle updater in the same EKF). Never imported by the product package.
"""
import numpy as np

from . import pyoracle as po


class OracleBackend:
    def __init__(self, threads=8, icp_normals="bins", icp_max_match_dist=None):
        # icp_*: variants of the restated libpointmatcher chain (icp_oracle.icp): the defaults are the reference's
        self.icp_normals, self.icp_max_match_dist = icp_normals, icp_max_match_dist
        self.map = np.zeros((0, 4), np.float32)
        self.init = False
        self.threads = threads
        self.last_T = None
        self.last_icp_success = True
        self.icp_runs = self.icp_ok = self.icp_iters = 0

    def close(self):
        pass

    def initialized(self):
        return self.init

    def _scan(self, scan, op, downsample):
        sc = po.range_filter(scan["xyz"], scan["t_rel"], op.max_range, op.min_range)  # LiDARData ctor
        if downsample and len(sc):
            sc, _ = po.voxel_downsample(sc, op.raw_downsample_size)                   # LidarHelper::downsample
        return sc

    def build_map_from_scan(self, scan, op, downsample=True):
        self.map = self._scan(scan, op, op.raw_do_downsample and downsample).copy()   # init_map_local -> Build
        self.init = len(self.map) > 0

    def scan_update(self, scan, op, T_LtoM, consts, P):
        from mins_b200 import capi
        sc = self._scan(scan, op, op.raw_do_downsample)
        if len(sc) == 0:
            return capi.ML_SOFT_NO_ROWS, None, None, capi.ml_counters()
        self.last_icp_success = True
        if op.map_use_icp:                                                            # LidarHelper.cpp:156-187
            T_LtoM, self.last_icp_success = self._icp(scan, sc, op, T_LtoM)
        self.last_T = np.array(T_LtoM, np.float32).reshape(4, 4)
        sm = po.transform_f32(sc, T_LtoM)                                             # transform_to_map
        ref = po.scan_linsys(consts, P, self.map, sc, sm, threads=self.threads, want_compress=True)
        if ref["cnt"].n_acc == 0:
            return capi.ML_SOFT_NO_ROWS, None, None, ref["cnt"]
        return 0, ref["Hc"], ref["rc"], ref["cnt"]

    def _icp(self, scan, sc, op, T_init):
        from . import icp_oracle
        r = icp_oracle.icp(sc[:, :3], self.map, T_init, op.map_icp_dist, seed=scan["key"] + 1000003 * scan.get("icp_tries", 0),
                           normals=self.icp_normals, max_match_dist=self.icp_max_match_dist)
        scan["icp_tries"] = scan.get("icp_tries", 0) + 1
        self.icp_runs += 1
        self.icp_ok += int(r["success"])
        self.icp_iters += r["iterations"]
        return r["T_LtoM"], bool(r["success"])

    def align(self, scan, op, T_init):
        """register_scan's retry of transform_to_map (LidarHelper.cpp:198)"""
        if not op.map_use_icp:
            return np.array(T_init, np.float32).reshape(4, 4), True
        return self._icp(scan, self._scan(scan, op, op.raw_do_downsample), op, T_init)

    def register_scan(self, scan, op, T_LtoM):
        sc = po.range_filter(scan["xyz"], scan["t_rel"], op.max_range, op.min_range)  # pointcloud_original
        sm = po.transform_f32(sc, T_LtoM)
        self.map = po.map_insert(self.map, sm, op.map_do_downsample, op.map_downsample_size)
        return len(self.map)

    def reanchor(self, T_old_to_new, now_minus_FT, op):
        self.map = po.reanchor(self.map, T_old_to_new, now_minus_FT, op.map_decay_time, op.map_decay_dist, op.map_do_downsample,
                               op.map_downsample_size)
        return len(self.map)

    def map_size(self):
        return len(self.map)
