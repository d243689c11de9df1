"""The simulator's LiDAR model: planes from a 2-D floor plan and ray/plane intersection (numpy, host side).

Restates mins/src/sim/Simulator.cpp:990-1096 (load_plane_data: every wall segment becomes a quad z in [-1, 4], plus floor
and ceiling over the bounding box), mins/src/sim/SimulationPlane.h:43-95 (plane equation from three corners, in-quad test
with the four corner dot products) and Simulator.cpp:742-815 (get_lidar_pointcloud: ray directions from v/h angles, nearest
hit, range gate, fp32 noise). The CUDA ray-caster (ml_sim_raycast) is checked against this file.
"""
import numpy as np


class PlaneWorld:
    def __init__(self, segments_m, z_lo=-1.0, z_hi=4.0):
        """segments_m: (S, 4) wall segments [x0, y0, x1, y1] in metres."""
        seg = np.asarray(segments_m, np.float64).reshape(-1, 4)
        quads = []
        for x0, y0, x1, y1 in seg:  # Simulator.cpp:1053-1067: tl, tr, bl, br
            quads.append([[x0, y0, z_hi], [x1, y1, z_hi], [x0, y0, z_lo], [x1, y1, z_lo]])
        mnx, mxx = min(seg[:, 0].min(), seg[:, 2].min()), max(seg[:, 0].max(), seg[:, 2].max())
        mny, mxy = min(seg[:, 1].min(), seg[:, 3].min()), max(seg[:, 1].max(), seg[:, 3].max())
        for z in (z_lo, z_hi):  # :1078-1090 floor and ceiling
            quads.append([[mnx, mny, z], [mnx, mxy, z], [mxx, mny, z], [mxx, mxy, z]])
        q = np.array(quads)  # (P, 4, 3): tl, tr, bl, br
        self.quads = q
        tl, tr, bl, br = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        N = np.cross(tr - tl, bl - tl)  # SimulationPlane.h:46-53
        self.abcd = np.concatenate([N, -(N * tl).sum(1, keepdims=True)], 1)
        unit = lambda v: v / np.linalg.norm(v, axis=1, keepdims=True)
        self.V1, self.V2, self.V3, self.V4 = unit(tr - tl), unit(bl - tl), unit(tr - br), unit(bl - br)
        self.tl, self.br = tl, br
        self.bounds = (mnx, mxx, mny, mxy)

    def packed(self):
        """(P, 22) float64 rows [A B C D | tl | br | V1 | V2 | V3 | V4]: the layout the CUDA ray-caster takes."""
        return np.ascontiguousarray(np.concatenate([self.abcd, self.tl, self.br, self.V1, self.V2, self.V3, self.V4], 1))

    def raycast(self, origin, dirs):
        """nearest positive hit distance per ray (inf if none). origin (3,), dirs (N,3) in the world frame."""
        o = np.asarray(origin, np.float64)
        d = np.asarray(dirs, np.float64)
        best = np.full(len(d), np.inf)
        A = self.abcd
        num = -(A[:, :3] @ o + A[:, 3])  # (P,)
        for s in range(0, len(d), 8192):
            dd = d[s:s + 8192]
            with np.errstate(divide="ignore", invalid="ignore"):
                t = num[None, :] / (dd @ A[:, :3].T)  # (n, P)  SimulationPlane.h:76
            pt = o[None, None, :] + t[:, :, None] * dd[:, None, :]
            U1 = pt - self.tl[None]
            U2 = pt - self.br[None]
            with np.errstate(divide="ignore", invalid="ignore"):
                U1 = U1 / np.linalg.norm(U1, axis=2, keepdims=True)
                U2 = U2 / np.linalg.norm(U2, axis=2, keepdims=True)
            ok = (t > 0) & ((U1 * self.V1[None]).sum(2) > 0) & ((U1 * self.V2[None]).sum(2) > 0) & \
                 ((U2 * self.V3[None]).sum(2) > 0) & ((U2 * self.V4[None]).sum(2) > 0)  # :94
            best[s:s + 8192] = np.where(ok, t, np.inf).min(1)
        return best


def load_floorplan(path):
    """Simulator::load_plane_data (mins/src/sim/Simulator.cpp:990-1096): lines starting with '#' are comments; a line with ONE
    number sets the units-to-metres factor; a line with FOUR comma-separated numbers is a wall segment
    (start_x, start_y, end_x, end_y) in file units. Returns (S, 4) segments in metres (PlaneWorld extrudes them to z in [-1, 4]
    and adds floor and ceiling over their bounding box, :1053-1090)."""
    units, segs = 1.0, []
    with open(path) as f:
        for line in f:
            if line.find("#") == 0:  # `!current_line.find("#")`: only lines that START with the comment sign
                continue
            data = [float(x) for x in line.strip().split(",") if x.strip() != ""]
            if len(data) == 1:
                units = data[0]
            elif len(data) == 4:
                segs.append(data)
    if not segs:
        raise ValueError(f"no plane in {path}")
    return np.asarray(segs, np.float64) * units


def synthetic_floorplan(seed=0, half_x=20.0, half_y=12.5):
    """An indoor hall with rooms along the walls and free-standing pillars (wall segments in metres), in the spirit of
    mins_data/FloorPlans/UD_Warehouse.txt (outer rectangle + rectangular obstacles)."""
    rng = np.random.default_rng(seed + 1000)
    segs = []

    def rect(x0, y0, x1, y1):
        segs.extend([[x0, y0, x0, y1], [x0, y0, x1, y0], [x0, y1, x1, y1], [x1, y0, x1, y1]])

    rect(-half_x, -half_y, half_x, half_y)
    x = -half_x
    while x < half_x - 3:  # shelves / rooms along the two long walls
        w = rng.uniform(2.0, 5.0)
        d = rng.uniform(0.8, 1.6)
        rect(x, -half_y, min(x + w, half_x), -half_y + d)
        rect(x + rng.uniform(0, 1), half_y - d, min(x + w, half_x), half_y)
        x += w + rng.uniform(0.5, 2.0)
    for _ in range(10):  # pillars and crates away from the centre line of the loop
        cx, cy = rng.uniform(-half_x + 2, half_x - 2), rng.uniform(-half_y + 2, half_y - 2)
        s = rng.uniform(0.4, 1.2)
        rect(cx - s, cy - s, cx + s, cy + s)
    return np.array(segs)


def lidar_ray_dirs(v_angles_deg, h_start=-180.0, h_end=180.0, h_res=2.0):
    """ray bearings in the LiDAR frame, v-major like Simulator.cpp:756-757 (for v: for h), with
    d = [sin(ph) cos(th), sin(ph) sin(th), cos(ph)], ph = 90 - v, th = h (:767-773). h_angles as OptionsLidar.cpp:93-98."""
    h = np.arange(h_start, h_end + 1e-9, h_res)
    v = np.asarray(v_angles_deg, np.float64)
    V, Hh = np.meshgrid(v, h, indexing="ij")
    ph, th = np.deg2rad(90.0 - V).ravel(), np.deg2rad(Hh).ravel()
    return np.stack([np.sin(ph) * np.cos(th), np.sin(ph) * np.sin(th), np.cos(ph)], 1)


def simulate_scan(world, R_GtoI, p_IinG, R_ItoL, p_IinL, dirs_L, min_range, max_range, noise, rng):
    """Simulator::get_lidar_pointcloud (Simulator.cpp:742-815): returns (n, 3) float32 points in the LiDAR frame."""
    p_LinI = -R_ItoL.T @ p_IinL
    origin = p_IinG + R_GtoI.T @ p_LinI
    dirs_G = dirs_L @ (R_GtoI.T @ R_ItoL.T).T  # R_ItoG R_LtoI d
    d = world.raycast(origin, dirs_G)
    keep = ~((d > max_range) | (d < min_range))
    pts = (dirs_L[keep] * d[keep, None]).astype(np.float32)
    if noise > 0:
        pts = pts + (np.float32(noise) * rng.standard_normal(pts.shape).astype(np.float32))
    return pts
