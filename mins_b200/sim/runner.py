"""Closed-loop IMU + LiDAR simulation (BASELINE.json configs 0 and 4): the run_simulation loop of
mins/src/run_simulation.cpp:71-130 restated — B-spline ground truth, IMU at 200 Hz, LiDAR at 10 Hz ray-cast against a floor
plan, clone window at 20 Hz / 1 s, LiDAR updater behind a backend (GPU C ABI or CPU oracle) — and the evaluator's ATE
(mins_eval/src/functions/ResultTrajectory.cpp:70-97: se3 alignment, per-pose position norm and so3-log angle, RMSE).
"""
import os
import time
from dataclasses import dataclass, field

import numpy as np

from ..lidar import OptionsLidar
from .ekf import State
from .se3 import exp_so3, log_so3, rot_to_quat_jpl
from .traj import BsplineSE3, load_bspline_dataset, synthetic_control_poses
from .updater import LidarUpdater
from .world import PlaneWorld, lidar_ray_dirs, load_floorplan, simulate_scan, synthetic_floorplan

# the MINS simulation's own inputs (mins_data/BSplineDataset/UD_Warehouse.txt, mins_data/FloorPlans/UD_Warehouse.txt: the
# dataset the reference's CI evaluates, .github/workflows/evaluation.yml), committed as fixtures
_GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tests", "golden")
UD_WAREHOUSE = (os.path.join(_GOLDEN, "UD_Warehouse_bspline.txt"), os.path.join(_GOLDEN, "UD_Warehouse_floorplan.txt"))


@dataclass
class SimConfig:
    seed: int = 0
    duration: float = 20.0           # seconds of estimator time
    freq_imu: float = 200.0          # config_simulation.yaml: freq_imu
    freq_lidar: float = 10.0         # freq_lidar
    clone_freq: float = 20.0         # config_estimator.yaml: clone_freq
    window_size: float = 1.0         # window_size
    intr_order: int = 3
    gyro_noise: float = 1.6968e-3    # config_imu.yaml
    accel_noise: float = 2.0e-2
    gyro_walk: float = 1.9393e-5
    accel_walk: float = 3.0e-3
    v_angles: tuple = (-80, -70, -60, -50, -40, -30, -20, -10, 0, 10, 20, 30, 40, 50, 60, 70, 80)  # config_lidar.yaml:38
    h_resolution: float = 2.0
    world_seed: int = 0              # the floor plan and trajectory are shared by all Monte-Carlo seeds
    init_ori_err: float = 0.01       # rad, 1 sigma of the initial perturbation
    init_pos_err: float = 0.02       # m
    use_lidar: bool = True
    scan_source: str = "numpy"       # "numpy": host ray-caster (world.py); "cuda": ml_sim_raycast (Philox noise keyed by seed, scan index)
    bspline_path: str = ""           # BSplineDataset file ("" = the synthetic loop); UD_WAREHOUSE[0] for the reference's
    floorplan_path: str = ""         # FloorPlans file ("" = the synthetic hall); UD_WAREHOUSE[1]
    remove_noise: bool = False       # config_simulation.yaml remove_noise: noise-free IMU, scans and initial state (P0 unchanged)
    use_gt_lidar: bool = False       # config_init.yaml use_gt_lidar: seed the map from a dense scan at the true pose
                                     # (Initializer::init_lidar_sim, mins/src/init/Initializer.cpp:308-340)
    op: OptionsLidar = field(default_factory=lambda: OptionsLidar(raw_do_downsample=True, map_use_icp=False))


def se3_align(est_p, gt_p):
    """ov_eval AlignTrajectory se3 (Umeyama without scale): R, t with gt ~ R est + t."""
    mu_e, mu_g = est_p.mean(0), gt_p.mean(0)
    W = (gt_p - mu_g).T @ (est_p - mu_e)
    U, _, Vt = np.linalg.svd(W)
    S = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        S[2, 2] = -1
    R = U @ S @ Vt
    return R, mu_g - R @ mu_e


def ate(est, gt):
    """est, gt: lists of (t, R_GtoI, p). Returns (rmse_ori_deg, rmse_pos_m) after se3 alignment (ResultTrajectory.cpp:70-97)."""
    ep = np.array([e[2] for e in est])
    gp = np.array([g[2] for g in gt])
    R, t = se3_align(ep, gp)
    epa = ep @ R.T + t
    e_pos = np.linalg.norm(epa - gp, axis=1)
    e_ori = []
    for e, g in zip(est, gt):
        R_est_ItoG = R @ e[1].T
        e_ori.append(np.linalg.norm(log_so3(g[1] @ R_est_ItoG)))
    e_ori = np.rad2deg(np.array(e_ori))
    return float(np.sqrt(np.mean(e_ori**2))), float(np.sqrt(np.mean(e_pos**2)))


def run(cfg, backend):
    """One simulation run. Returns a dict with the estimated and ground-truth trajectories, ATE, NEES and timing."""
    rng_imu = np.random.default_rng([cfg.seed, 1])
    rng_ldr = np.random.default_rng([cfg.seed, 2])
    rng_init = np.random.default_rng([cfg.seed, 3])
    spline = BsplineSE3(load_bspline_dataset(cfg.bspline_path) if cfg.bspline_path else synthetic_control_poses(cfg.world_seed, cfg.duration + 6.0))
    world = PlaneWorld(load_floorplan(cfg.floorplan_path) if cfg.floorplan_path else synthetic_floorplan(cfg.world_seed))
    op = cfg.op
    dirs_L = lidar_ray_dirs(cfg.v_angles, -180.0, 180.0, cfg.h_resolution)
    # lidar0 extrinsics of config_lidar.yaml:33-37: T_imu_lidar = [R_LtoI | p_LinI]
    R_LtoI = np.array([[1.0, 0, 0], [0, 0.9659258, -0.2588190], [0, 0.2588190, 0.9659258]])
    u, _, vt = np.linalg.svd(R_LtoI)
    R_ItoL = (u @ vt).T
    p_IinL = -R_ItoL @ np.array([0.1, 0.1, 0.1])

    dt_imu = 1.0 / cfg.freq_imu
    t0 = np.ceil((spline.t_start + 0.1) * cfg.clone_freq) / cfg.clone_freq
    k0 = spline.get_kinematics(t0)
    nz = 0.0 if cfg.remove_noise else 1.0
    R0 = exp_so3(nz * rng_init.normal(0, cfg.init_ori_err, 3)) @ k0[0]
    p0 = k0[1] + nz * rng_init.normal(0, cfg.init_pos_err, 3)
    P0 = np.diag(np.r_[np.full(3, cfg.init_ori_err**2), np.full(3, cfg.init_pos_err**2), np.full(3, 0.01**2), np.full(3, 1e-4**2 + 1e-8),
                       np.full(3, 0.02**2)])
    st = State(R0, p0, k0[3].copy(), np.zeros(3), np.zeros(3), P0, intr_order=cfg.intr_order, R_ItoL=R_ItoL, p_IinL=p_IinL)
    st.time = t0
    upd = LidarUpdater(st, op, backend)
    noises = (cfg.gyro_noise, cfg.accel_noise, cfg.gyro_walk, cfg.accel_walk)
    bg_true, ba_true = np.zeros(3), np.zeros(3)

    def imu_meas(t):
        nonlocal bg_true, ba_true
        R, _, w, _, a = spline.get_kinematics(t)
        sd = 1.0 / np.sqrt(dt_imu)
        wm = w + bg_true + nz * cfg.gyro_noise * sd * rng_imu.standard_normal(3)
        am = R @ (a + st.g) + ba_true + nz * cfg.accel_noise * sd * rng_imu.standard_normal(3)
        bg_true = bg_true + nz * cfg.gyro_walk * np.sqrt(dt_imu) * rng_imu.standard_normal(3)
        ba_true = ba_true + nz * cfg.accel_walk * np.sqrt(dt_imu) * rng_imu.standard_normal(3)
        return wm, am

    caster = None
    if cfg.use_lidar and cfg.scan_source == "cuda":
        from ..lidar import LidarMap
        caster = LidarMap(max_scan_points=max(1 << 17, len(dirs_L)), max_map_points=1024, device=getattr(backend, "device", 0))
        caster.sim_set_world(world.packed())
        caster.sim_set_rays(dirs_L)
    p_LinI = -R_ItoL.T @ p_IinL
    if cfg.use_lidar and cfg.use_gt_lidar:
        # Initializer::init_lidar_sim: v_angles = -90..89; the 180 values -90..89 are APPENDED to the h_angles the options already
        # hold (-180:h_res:180): 180 x 361 rays for the simulation YAML; the scan is taken at the true pose with the usual noise
        # and becomes the map as it is (LiDARData ctor + init_map_local: range filter, no down-sampling)
        h_ang = np.concatenate([np.arange(-180.0, 180.0 + 1e-9, cfg.h_resolution), np.arange(-90.0, 90.0)])
        V, Hh = np.meshgrid(np.arange(-90.0, 90.0), h_ang, indexing="ij")
        ph, th = np.deg2rad(90.0 - V).ravel(), np.deg2rad(Hh).ravel()
        dense = np.stack([np.sin(ph) * np.cos(th), np.sin(ph) * np.sin(th), np.cos(ph)], 1)
        if caster is not None:  # same generator as the C++ harness: Philox keyed by (seed, 0xffffffff)
            caster.sim_set_rays(dense)
            xyz0 = caster.sim_raycast(k0[1] + k0[0].T @ p_LinI, k0[0].T @ R_ItoL.T, op.min_range, op.max_range, noise=nz * op.raw_noise,
                                      seed=(int(cfg.seed) << 32) | 0xffffffff)[:, :3]
            caster.sim_set_rays(dirs_L)
        else:
            xyz0 = simulate_scan(world, k0[0], k0[1], R_ItoL, p_IinL, dense, op.min_range, op.max_range, nz * op.raw_noise, rng_ldr)
        upd.init_map_gt(t0, xyz0)
    n_steps = int(round(min(cfg.duration, spline.t_end - 0.05 - t0) * cfg.freq_imu))
    imu_per_clone = int(round(cfg.freq_imu / cfg.clone_freq))
    imu_per_scan = int(round(cfg.freq_imu / cfg.freq_lidar))
    est, gt, nees_p = [], [], []
    t_ldr = 0.0
    prev = imu_meas(t0)
    st.augment_clone()
    for k in range(1, n_steps + 1):
        t = t0 + k * dt_imu
        cur = imu_meas(t)
        st.propagate(0.5 * (prev[0] + cur[0]), 0.5 * (prev[1] + cur[1]), dt_imu, noises)
        st.time = t  # keep clone keys on the exact grid
        prev = cur
        if k % imu_per_clone == 0:  # SystemManager::feed_measurement_imu: clone, marginalise, re-anchor the map (:99-105)
            st.augment_clone()
            while st.newest_clone_time() - st.oldest_clone_time() > cfg.window_size + 1e-9:
                st.marginalize_oldest()
            if cfg.use_lidar:
                upd.propagate_map_frame()
            R_gt, p_gt = spline.get_pose(t)
            est.append((t, st.R.copy(), st.p.copy()))
            gt.append((t, R_gt, p_gt))
            e = st.p - p_gt
            nees_p.append(float(e @ np.linalg.solve(st.P[3:6, 3:6], e)))
        if cfg.use_lidar and k % imu_per_scan == 0:  # SystemManager::feed_measurement_lidar (:177-185)
            R_gt, p_gt = spline.get_pose(t)
            if caster is not None:
                xyz = caster.sim_raycast(p_gt + R_gt.T @ p_LinI, R_gt.T @ R_ItoL.T, op.min_range, op.max_range, noise=nz * op.raw_noise,
                                         seed=(int(cfg.seed) << 32) | (k // imu_per_scan))[:, :3]
            else:
                xyz = simulate_scan(world, R_gt, p_gt, R_ItoL, p_IinL, dirs_L, op.min_range, op.max_range, nz * op.raw_noise, rng_ldr)
            c0 = time.perf_counter()
            upd.feed_measurement(t, xyz)
            upd.try_update()
            t_ldr += time.perf_counter() - c0
    if caster is not None:
        caster.close()
    rmse_ori, rmse_pos = ate(est, gt)
    traj = np.array([[e[0], *e[2], *rot_to_quat_jpl(e[1])] for e in est])
    return dict(rmse_ori_deg=rmse_ori, rmse_pos_m=rmse_pos, nees_pos=float(np.mean(nees_p)) / 3.0, n_updates=upd.n_updates,
                n_scans=n_steps // imu_per_scan, ldr_wall_s=t_ldr, map_points=backend.map_size() if cfg.use_lidar else 0,
                traj=traj, final_pos_err=float(np.linalg.norm(est[-1][2] - gt[-1][2])),
                n_acc=int(sum(c.n_acc for c in upd.counters)), n_points=int(sum(c.n_points for c in upd.counters)),
                icp_runs=getattr(backend, "icp_runs", 0), icp_ok=getattr(backend, "icp_ok", 0), icp_iters=getattr(backend, "icp_iters", 0))
