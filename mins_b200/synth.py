"""Synthetic LiDAR workloads of the shapes BASELINE.json names: a street-canyon-like world of a
ground plane and axis-aligned boxes, a spinning multi-beam LiDAR ray-cast into it (restating the simulator's
ray/plane model, mins/src/sim/Simulator.cpp:742-815), a voxel-spaced local map of the same world expressed in the
map (anchor LiDAR) frame, and the per-scan constants UpdaterLidar::update would derive from the EKF state.
numpy only; deterministic per seed. Used by tests/ and bench.py for inputs — not part of the device path.
"""
import numpy as np

from .lidar import OptionsLidar, make_scan_consts


def rot_zyx(yaw, pitch, roll):
    cy, sy, cp, sp, cr, sr = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch), np.cos(roll), np.sin(roll)
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1.0]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    return Rz @ Ry @ Rx


def exp_so3(w):
    th = np.linalg.norm(w)
    K = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    if th < 1e-12:
        return np.eye(3) + K
    return np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / th**2 * K @ K


class World:
    """ground plane z = zg over [-ext, ext]^2 plus boxes (xmin, xmax, ymin, ymax, height above ground)."""

    def __init__(self, seed, ext, n_boxes, zg=-1.9):
        rng = np.random.default_rng(seed)
        self.ext, self.zg = float(ext), float(zg)
        boxes = []
        # two facades 12 m apart along x (the street), broken into blocks
        for side in (-1, 1):
            x = -ext
            while x < ext:
                L = rng.uniform(15, 40)
                depth = rng.uniform(8, 20)
                h = rng.uniform(6, 25)
                y0 = side * 6.0
                boxes.append((x, min(x + L, ext), min(y0, y0 + side * depth), max(y0, y0 + side * depth), h))
                x += L + rng.uniform(0, 6)
        # free-standing boxes ("vehicles", kiosks, far buildings)
        for _ in range(n_boxes):
            cx, cy = rng.uniform(-ext, ext, 2)
            if abs(cy) < 30 and abs(cy) > 4.5:
                cy = np.sign(cy) * rng.uniform(30, ext)
            if abs(cy) <= 4.5:
                sx, sy, h = rng.uniform(3, 5), rng.uniform(1.6, 2.0), rng.uniform(1.4, 2.5)
                cy = np.sign(cy + 1e-9) * rng.uniform(2.5, 4.5)
            else:
                sx, sy, h = rng.uniform(8, 30), rng.uniform(8, 30), rng.uniform(4, 30)
            boxes.append((cx - sx / 2, cx + sx / 2, cy - sy / 2, cy + sy / 2, h))
        self.boxes = np.array(boxes, dtype=np.float64)

    def raycast(self, origin, dirs, max_range):
        """nearest hit distance per ray (inf if none). dirs: (N,3) unit vectors in the world frame."""
        o = np.asarray(origin, np.float64)
        d = np.asarray(dirs, np.float64)
        N = len(d)
        t_best = np.full(N, np.inf)
        with np.errstate(divide="ignore", invalid="ignore"):
            tg = (self.zg - o[2]) / d[:, 2]
            hit = (tg > 0) & np.isfinite(tg)
            px, py = o[0] + tg * d[:, 0], o[1] + tg * d[:, 1]
            hit &= (np.abs(px) <= self.ext) & (np.abs(py) <= self.ext)
            t_best[hit] = tg[hit]
            inv = 1.0 / d
            lo = np.array([self.boxes[:, 0], self.boxes[:, 2], np.full(len(self.boxes), self.zg)]).T
            hi = np.array([self.boxes[:, 1], self.boxes[:, 3], self.zg + self.boxes[:, 4]]).T
            for s in range(0, N, 16384):
                e = min(N, s + 16384)
                t1 = (lo[None, :, :] - o[None, None, :]) * inv[s:e, None, :]
                t2 = (hi[None, :, :] - o[None, None, :]) * inv[s:e, None, :]
                tmin = np.nanmax(np.minimum(t1, t2), axis=2)
                tmax = np.nanmin(np.maximum(t1, t2), axis=2)
                ok = (tmax >= np.maximum(tmin, 0.0)) & (tmin > 0)
                tb = np.where(ok, tmin, np.inf).min(axis=1)
                t_best[s:e] = np.minimum(t_best[s:e], tb)
        t_best[t_best > max_range] = np.inf
        return t_best

    def sample_surfaces(self, spacing, rng, centre, radius):
        """one jittered point per spacing x spacing surface patch within `radius` of `centre` (world frame)."""
        pts = []

        def patch(u0, u1, v0, v1):
            nu, nv = max(1, int((u1 - u0) / spacing)), max(1, int((v1 - v0) / spacing))
            u = u0 + (np.arange(nu) + 0.5) * (u1 - u0) / nu
            v = v0 + (np.arange(nv) + 0.5) * (v1 - v0) / nv
            U, V = np.meshgrid(u, v, indexing="ij")
            U = U + rng.uniform(-0.3, 0.3, U.shape) * spacing
            V = V + rng.uniform(-0.3, 0.3, V.shape) * spacing
            return U.ravel(), V.ravel()

        cx, cy = centre[0], centre[1]
        gx0, gx1 = max(-self.ext, cx - radius), min(self.ext, cx + radius)
        gy0, gy1 = max(-self.ext, cy - radius), min(self.ext, cy + radius)
        U, V = patch(gx0, gx1, gy0, gy1)
        inside = np.zeros(len(U), bool)
        for b in self.boxes:
            inside |= (U > b[0]) & (U < b[1]) & (V > b[2]) & (V < b[3])
        pts.append(np.stack([U, V, np.full_like(U, self.zg)], 1)[~inside])
        for b in self.boxes:
            x0, x1, y0, y1, h = b
            z0, z1 = self.zg, self.zg + h
            for (xa, xb, ya, yb) in ((x0, x1, y0, y0), (x0, x1, y1, y1), (x0, x0, y0, y1), (x1, x1, y0, y1)):
                if xa == xb:
                    U, V = patch(ya, yb, z0, z1)
                    pts.append(np.stack([np.full_like(U, xa), U, V], 1))
                else:
                    U, V = patch(xa, xb, z0, z1)
                    pts.append(np.stack([U, np.full_like(U, ya), V], 1))
            U, V = patch(x0, x1, y0, y1)
            pts.append(np.stack([U, V, np.full_like(U, z1)], 1))
        P = np.concatenate(pts, 0)
        keep = np.linalg.norm(P - np.asarray(centre)[None, :], axis=1) <= radius
        return P[keep]


def beam_dirs(v_angles_deg, n_azimuth):
    """unit ray directions in the LiDAR frame, azimuth-major like Simulator.cpp:756-762 (for h: for v)."""
    az = np.deg2rad(-180.0 + 360.0 * np.arange(n_azimuth) / n_azimuth)
    el = np.deg2rad(np.asarray(v_angles_deg, np.float64))
    A, E = np.meshgrid(az, el, indexing="ij")
    return np.stack([np.cos(E) * np.cos(A), np.cos(E) * np.sin(A), np.sin(E)], -1).reshape(-1, 3)


SHAPES = {
    # name: (beam elevation angles, azimuth steps, sensors, map points, world half-extent, free boxes, map radius)
    "G-64": (np.linspace(-24.8, 2.0, 64), 2048, 1, 1_000_000, 160.0, 60, 150.0),
    "G-vlp": (np.arange(-15.0, 15.1, 2.0), 1000, 2, 200_000, 80.0, 40, 70.0),
    "G-small": (np.arange(-15.0, 15.1, 2.0), 128, 1, 20_000, 30.0, 8, 28.0),
}


def lagrange_weights(ts, t):
    w = np.ones(len(ts))
    for j in range(len(ts)):
        for m in range(len(ts)):
            if m != j:
                w[j] *= (t - ts[m]) / (ts[j] - ts[m])
    return w


def make_workload(shape="G-64", seed=None, op=None, variant="plain", map_points=None, noise=0.02, map_radius=None):
    """Returns a dict: raw scan (n,4 float32, PointXYZ layout with pad), map (M,4 float32 in the map frame),
    options, scan constants, T_LtoM, P, and the scan time stamp.

    variant: "plain" (total_hx 48), "overlap" (I and A share clones, 30), "calib" (54), "calib_dt" (55),
             "at_clone" (both times are clones: no interpolation-error inflation).
    """
    v_angles, n_az, n_sensors, M_target, ext, n_boxes, map_radius_ = SHAPES[shape]
    map_radius = map_radius_ if map_radius is None else float(map_radius)  # e.g. map_decay_dist of the option set
    if map_points is not None:
        M_target = int(map_points)
    if seed is None:
        seed = {"G-64": 64, "G-vlp": 38}.get(shape, 7)
    rng = np.random.default_rng(seed)
    op = op or OptionsLidar(raw_do_downsample=False)
    world = World(seed, ext, n_boxes)

    # calibration: lidar0 of mins/config/simulation/config_lidar.yaml:33-37 (T_imu_lidar = [R_LtoI, p_LinI])
    RLtoI = np.array([[1.0, 0, 0], [0, 0.9659258, -0.2588190], [0, 0.2588190, 0.9659258]])
    u, _, vt = np.linalg.svd(RLtoI)
    RLtoI = u @ vt
    RItoL = RLtoI.T
    pLinI = np.array([0.1, 0.1, 0.1])
    pIinL = -RItoL @ pLinI

    # true IMU poses: anchor A and scan time I (10 m/s along the street, slight turn)
    RGtoA = rot_zyx(0.05, 0.01, -0.26).T
    pAinG = np.array([0.0, 0.5, 0.0])
    RGtoI = (rot_zyx(0.05, 0.01, -0.26) @ exp_so3(np.array([0.004, -0.006, 0.03]))).T
    pIinG = pAinG + np.array([1.0, 0.03, 0.01])

    def lidar_pose(RGtoX, pXinG):
        return RItoL @ RGtoX, pXinG + RGtoX.T @ pLinI  # R_GtoL, p_LinG

    RGtoM, pMinG = lidar_pose(RGtoA, pAinG)
    RGtoL, pLinG = lidar_pose(RGtoI, pIinG)

    # ---- map: surface samples at the map voxel spacing, in the map frame, nearest M_target to the map origin ----
    Pw = world.sample_surfaces(op.map_downsample_size, rng, pMinG, map_radius)
    Pw = Pw + rng.normal(0, noise, Pw.shape)
    d = np.linalg.norm(Pw - pMinG[None, :], axis=1)
    if len(Pw) > M_target:
        sel = np.argpartition(d, M_target)[:M_target]
        Pw = Pw[np.sort(sel)]
    Pm = (Pw - pMinG[None, :]) @ RGtoM.T
    map_xyzi = np.zeros((len(Pm), 4), np.float32)
    map_xyzi[:, :3] = Pm.astype(np.float32)
    map_xyzi[:, 3] = rng.uniform(0.0, 5.0, len(Pm)).astype(np.float32)
    map_xyzi = map_xyzi[rng.permutation(len(map_xyzi))]

    # ---- scan: ray-cast from the scan-time LiDAR pose(s); rays without a return are (0,0,0) ----
    dirs_L = beam_dirs(v_angles, n_az)
    scans = []
    for s in range(n_sensors):
        Rs = rot_zyx(0.0, 0.0, 0.0) if s == 0 else rot_zyx(np.pi / 2, 0.35, 0.0)
        dl = dirs_L @ Rs.T
        dw = dl @ RGtoL  # world direction = R_LtoG * d_L
        t = world.raycast(pLinG, dw, op.max_range * 1.2)
        p = dl * np.where(np.isfinite(t), t, 0.0)[:, None]
        p = p.astype(np.float32) + np.where(np.isfinite(t)[:, None], rng.normal(0, noise, p.shape), 0.0).astype(np.float32)
        scans.append(p)
    pts = np.concatenate(scans, 0)
    raw = np.zeros((len(pts), 4), np.float32)
    raw[:, :3] = pts
    raw[:, 3] = 1.0  # pcl::PointXYZ padding

    # ---- estimates: truth perturbed by 0.5 deg / 0.03 m; fej slightly different from est ----
    def perturb(R, p, so, sp):
        return exp_so3(rng.normal(0, so, 3)) @ R, p + rng.normal(0, sp, 3)

    so, sp = np.deg2rad(0.5) / np.sqrt(3) * 0.3, 0.03 / np.sqrt(3) * 0.3
    RGtoI_est, pIinG_est = perturb(RGtoI, pIinG, so, sp)
    RGtoA_est, pAinG_est = perturb(RGtoA, pAinG, so * 0.3, sp * 0.3)
    RGtoI_fej, pIinG_fej = perturb(RGtoI_est, pIinG_est, 1e-3, 2e-3)
    RGtoA_fej, pAinG_fej = perturb(RGtoA_est, pAinG_est, 1e-3, 2e-3)

    # ---- Jacobian blocks of the two interpolated poses (State.cpp:745-816 shapes) ----
    ts = np.array([0.0, 0.05, 0.10, 0.15])
    wI = lagrange_weights(ts, 0.05 + 0.05 * rng.uniform(0.2, 0.8))
    wA = lagrange_weights(ts, 0.05 + 0.05 * rng.uniform(0.2, 0.8))
    mk = lambda w: w * np.eye(6) + 0.02 * rng.normal(size=(6, 6))
    col0 = 0
    calib_col = -1
    extra_I, extra_A = [], []
    if variant in ("calib", "calib_dt"):
        calib_col = 0
        col0 = 6
        if variant == "calib_dt":
            extra_I = [(6, 0.3 * rng.normal(size=(6, 1)))]
            extra_A = [(6, 0.3 * rng.normal(size=(6, 1)))]
            col0 = 7
    colsI = [col0 + 6 * j for j in range(4)]
    colsA = [col0 + 6 * (j + 1) for j in range(4)] if variant == "overlap" else [col0 + 24 + 6 * j for j in range(4)]
    dTdxI = [(colsI[j], mk(wI[j])) for j in range(4)] + extra_I
    dTdxA = [(colsA[j], mk(wA[j])) for j in range(4)] + extra_A
    total_hx = max(colsA) + 6
    at_clone = variant == "at_clone"
    consts, T_LtoM = make_scan_consts(op, RGtoI_est, pIinG_est, RGtoA_est, pAinG_est, RGtoI_fej, pIinG_fej, RGtoA_fej,
                                      pAinG_fej, RItoL, pIinL, dTdxI, dTdxA, calib_col, total_hx, at_clone_m=at_clone,
                                      at_clone_a=at_clone, ie_o=2.5e-7, ie_p=1.0e-5)
    rp = np.random.default_rng(7)
    L = rp.normal(size=(total_hx, total_hx))
    P = 1e-4 * (L @ L.T / total_hx + np.eye(total_hx))
    return dict(shape=shape, seed=seed, raw=raw, map=map_xyzi, op=op, consts=consts, T_LtoM=T_LtoM, P=P, t_rel=5.0,
                total_hx=total_hx, n_rays=len(raw))
