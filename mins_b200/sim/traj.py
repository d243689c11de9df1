"""Ground-truth trajectory for the closed-loop harness: cumulative cubic B-spline on SE(3), restating
thirdparty/open_vins/ov_core/src/sim/BsplineSE3.cpp (feed_trajectory :24-100, get_pose :102-136), which the MINS simulator
uses (mins/src/sim/Simulator.cpp get_imu_pose / get_imu_velocity / get_imu_acceleration). Velocities and accelerations are
taken by 4th-order central differences of the pose (the reference differentiates its velocities the same way,
Simulator.cpp:985)."""
import numpy as np

from .se3 import exp_se3, inv_se3, log_se3, log_so3, quat_jpl_to_rot


class BsplineSE3:
    def __init__(self, traj_points):
        """traj_points: (N, 8) rows [t, px, py, pz, qx, qy, qz, qw] with q = JPL q_GtoI (the BSplineDataset format)."""
        tp = np.asarray(traj_points, np.float64)
        # BsplineSE3.cpp:28-35: control-point spacing = mean spacing of the input, at least 0.05 s
        dt = float(np.mean(np.diff(tp[:, 0])))
        self.dt = max(dt, 0.05)
        poses = {}
        for row in tp[:-1]:  # :40 (the reference drops the last point)
            T = np.eye(4)
            T[:3, :3] = quat_jpl_to_rot(row[4:8]).T  # R_ItoG
            T[:3, 3] = row[1:4]
            poses[row[0]] = T
        times = np.array(sorted(poses))
        Ts = [poses[t] for t in times]
        # :62-88 uniform control points by linear SE(3) interpolation between the bounding input poses
        self.ctrl_t, self.ctrl_T = [], []
        t = times[0]
        while True:
            k = np.searchsorted(times, t, side="right") - 1
            if k < 0 or k + 1 >= len(times):
                break
            t0, t1 = times[k], times[k + 1]
            lam = (t - t0) / (t1 - t0)
            T = exp_se3(lam * log_se3(Ts[k + 1] @ inv_se3(Ts[k]))) @ Ts[k]
            self.ctrl_t.append(t)
            self.ctrl_T.append(T)
            t += self.dt
        self.ctrl_t = np.array(self.ctrl_t)
        self.t_start = times[0] + 2 * self.dt  # :91
        self.t_end = self.ctrl_t[-3] if len(self.ctrl_t) > 4 else self.ctrl_t[-1]
        # the relative twists between consecutive control points are reused by every evaluation
        self._rel = [log_se3(inv_se3(self.ctrl_T[i]) @ self.ctrl_T[i + 1]) for i in range(len(self.ctrl_T) - 1)]

    def pose_T(self, t):
        """4x4 T_ItoG at time t (BsplineSE3.cpp:102-136), None outside the spline."""
        i1 = int(np.searchsorted(self.ctrl_t, t, side="right") - 1)
        if i1 < 1 or i1 + 2 >= len(self.ctrl_t):
            return None
        u = (t - self.ctrl_t[i1]) / (self.ctrl_t[i1 + 1] - self.ctrl_t[i1])
        b0 = (5 + 3 * u - 3 * u * u + u**3) / 6.0
        b1 = (1 + 3 * u + 3 * u * u - 2 * u**3) / 6.0
        b2 = u**3 / 6.0
        A0 = exp_se3(b0 * self._rel[i1 - 1])
        A1 = exp_se3(b1 * self._rel[i1])
        A2 = exp_se3(b2 * self._rel[i1 + 1])
        return self.ctrl_T[i1 - 1] @ A0 @ A1 @ A2

    def get_pose(self, t):
        """(R_GtoI, p_IinG) or None."""
        T = self.pose_T(t)
        if T is None:
            return None
        return T[:3, :3].T, T[:3, 3].copy()

    def get_kinematics(self, t, h=2e-3):
        """(R_GtoI, p, w_IinI, v_IinG, a_IinG) by central differences; None near the ends."""
        Ts = [self.pose_T(t + k * h) for k in (-2, -1, 0, 1, 2)]
        if any(T is None for T in Ts):
            return None
        p = [T[:3, 3] for T in Ts]
        R = [T[:3, :3].T for T in Ts]  # R_GtoI
        v = (p[0] - 8 * p[1] + 8 * p[3] - p[4]) / (12 * h)
        a = (-p[0] + 16 * p[1] - 30 * p[2] + 16 * p[3] - p[4]) / (12 * h * h)
        # R_GtoI(t+h) = exp_so3(-w h) R_GtoI(t)  (JPL kinematics)
        w1 = -log_so3(R[3] @ R[1].T) / (2 * h)
        w2 = -log_so3(R[4] @ R[0].T) / (4 * h)
        w = (4 * w1 - w2) / 3.0  # Richardson: 4th order
        return R[2], p[2].copy(), w, v, a


def load_bspline_dataset(path):
    """mins_data/BSplineDataset/*.txt (ov_core DatasetReader::load_simulated_trajectory format): rows
    [t, px, py, pz, qx, qy, qz, qw], '#' comments. Returns (N, 8)."""
    rows = []
    with open(path) as f:
        for line in f:
            line = line.strip()
            if not line or line.startswith("#"):
                continue
            v = [float(x) for x in line.replace(",", " ").split()]
            if len(v) == 8:
                rows.append(v)
    return np.asarray(rows, np.float64)


def synthetic_control_poses(seed=0, duration=60.0, rate=2.0, half_x=14.0, half_y=8.0, z0=0.6):
    """A smooth loop inside a 2*half_x by 2*half_y hall (BSplineDataset row format), for boxes without mins_data.
    Lissajous-like path with gentle roll/pitch and heading following the velocity direction."""
    rng = np.random.default_rng(seed)
    ph = rng.uniform(0, 2 * np.pi, 4)
    n = int(duration * rate) + 8
    t = np.arange(n) / rate
    w0 = 2 * np.pi / 45.0
    x = half_x * 0.75 * np.sin(w0 * t + ph[0])
    y = half_y * 0.7 * np.sin(2 * w0 * t + ph[1])
    z = z0 + 0.25 * np.sin(1.7 * w0 * t + ph[2])
    vx = np.gradient(x, t)
    vy = np.gradient(y, t)
    yaw = np.unwrap(np.arctan2(vy, vx))
    roll = 0.08 * np.sin(2.3 * w0 * t + ph[3])
    pitch = 0.06 * np.cos(1.9 * w0 * t + ph[0])
    from .se3 import rot_to_quat_jpl
    rows = []
    for k in range(n):
        cy, sy, cp, sp, cr, sr = np.cos(yaw[k]), np.sin(yaw[k]), np.cos(pitch[k]), np.sin(pitch[k]), np.cos(roll[k]), np.sin(roll[k])
        Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1.0]])
        Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
        Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
        R_ItoG = Rz @ Ry @ Rx
        q = rot_to_quat_jpl(R_ItoG.T)
        rows.append([t[k], x[k], y[k], z[k], *q])
    return np.array(rows)
