"""Shared ICP test scene: a local map built from two dense simulated scans of the synthetic hall (map frame = LiDAR frame of
the first pose) and a sparser scan from a displaced pose, so the true T_LtoM is known exactly."""
import numpy as np


def make_scene(seed=0, noise=0.01, dense_res=1.0, scan_res=2.0):
    from mins_b200.sim.se3 import exp_so3
    from mins_b200.sim.world import PlaneWorld, lidar_ray_dirs, synthetic_floorplan
    rng = np.random.default_rng(seed)
    world = PlaneWorld(synthetic_floorplan(seed))
    # the map needs comparable angular spacing in both directions: bins of <= 7 points taken along one sparse ring are
    # collinear and their PCA normal is meaningless (true of the reference filter as well)
    dense = lidar_ray_dirs(np.arange(-80, 81, 2.0 * dense_res), -180, 180, dense_res)
    sparse = lidar_ray_dirs(np.arange(-80, 81, 10), -180, 180, scan_res)

    def scan(R_LtoG, p_LinG, dirs):
        d = world.raycast(p_LinG, dirs @ R_LtoG.T)
        keep = (d > 0.5) & (d < 60.0)
        pts = dirs[keep] * d[keep, None]
        return (pts + noise * rng.standard_normal(pts.shape)).astype(np.float32)

    # anchor (map) pose and a second mapping pose 1 m away; the new scan is 0.4 m / 6 deg from the anchor
    R_M, p_M = exp_so3(np.array([0.02, -0.03, 0.4])), np.array([1.0, -1.5, 0.9])
    R_2, p_2 = R_M @ exp_so3(np.array([0.0, 0.01, 0.15])), p_M + np.array([0.8, 0.5, 0.05])
    R_L, p_L = R_M @ exp_so3(np.array([0.01, -0.02, 0.10])), p_M + np.array([0.35, 0.2, -0.03])

    def to_map(pts, R, p):  # LiDAR frame (R, p in G) -> map frame
        return ((pts.astype(np.float64) @ R.T + p - p_M) @ R_M).astype(np.float32)

    m1 = scan(R_M, p_M, dense)
    m2 = to_map(scan(R_2, p_2, dense), R_2, p_2)
    map_xyz = np.concatenate([m1, m2])
    # one point per 0.1 m voxel (first come), like the map's own down-sampling
    key = np.floor(map_xyz / 0.1).astype(np.int64)
    _, first = np.unique(key, axis=0, return_index=True)
    map_xyz = map_xyz[np.sort(first)]
    reading = scan(R_L, p_L, sparse)
    T_true = np.eye(4)
    T_true[:3, :3] = R_M.T @ R_L
    T_true[:3, 3] = R_M.T @ (p_L - p_M)
    T_init = np.eye(4)
    T_init[:3, :3] = exp_so3(np.deg2rad([0.8, -0.6, 1.2])) @ T_true[:3, :3]
    T_init[:3, 3] = T_true[:3, 3] + np.array([0.05, -0.04, 0.03])
    map_xyzi = np.zeros((len(map_xyz), 4), np.float32)
    map_xyzi[:, :3] = map_xyz
    return dict(map=map_xyzi, reading=reading, T_true=T_true, T_init=T_init.astype(np.float32))


def pose_error(T, T_true):
    """(rotation error [deg], translation error [m])"""
    T, T_true = np.asarray(T, np.float64), np.asarray(T_true, np.float64)
    dR = T[:3, :3] @ T_true[:3, :3].T
    ang = np.degrees(np.arccos(np.clip((np.trace(dR) - 1) / 2, -1, 1)))
    return float(ang), float(np.linalg.norm(T[:3, 3] - T_true[:3, 3]))
