"""Minimal IMU + clone-window EKF core for the closed-loop harness (numpy, fp64) — SURVEY.md §8(f) N1.

This is NOT part of the LiDAR hot path: it stands in for the untouched MINS EKF core (mins/src/state/{State,StateHelper,
Propagator}) which cannot be compiled in this image (no Eigen/Boost/ROS), so that the GPU LiDAR updater and the CPU-oracle
updater can be compared in closed loop on trajectory RMSE. Restated pieces, with the reference lines they follow:
  * clone window + marginalisation                     SystemManager.cpp:265-394, StateHelper::marginalize
  * polynomial pose interpolation and its Jacobians    State.cpp:428-520 (build_polynomial_data), :601-643, :675-826
  * marginal covariance / EKF update                   StateHelper.cpp:227-254, :121-200
  * IMU propagation (discrete, first order)            Propagator.cpp (mean: R <- exp(-w dt) R, v, p; Phi, Qd)
Error-state convention: JPL, R_true = exp_so3(-dth) R_hat (ov_type::JPLQuat::update), positions additive.
Simplifications (documented in DESIGN.md): polynomial interpolation is used for the estimate as well as the Jacobian
(the reference's `use_imu_res: false` branch), interpolation-error variances are constants.
"""
import numpy as np

from .se3 import Jl_so3, exp_so3, log_so3, skew


class Clone:
    __slots__ = ("R", "p", "R_fej", "p_fej", "idx")

    def __init__(self, R, p, idx):
        self.R, self.p = R.copy(), p.copy()
        self.R_fej, self.p_fej = R.copy(), p.copy()
        self.idx = idx  # first column in the covariance


class State:
    """state vector: [IMU 15: th p v bg ba | calib 6 (optional) | clones 6 each, oldest first]"""

    def __init__(self, R_GtoI, p, v, bg, ba, P0, intr_order=3, gravity=9.81, R_ItoL=None, p_IinL=None, calib_cov=None,
                 intr_err_ori=1e-6, intr_err_pos=1e-5):
        self.R, self.p, self.v, self.bg, self.ba = R_GtoI.copy(), p.copy(), v.copy(), bg.copy(), ba.copy()
        self.time = 0.0
        self.intr_order = intr_order
        self.g = np.array([0.0, 0.0, gravity])
        self.clones = {}  # time -> Clone (insertion order = time order)
        self.R_ItoL = np.eye(3) if R_ItoL is None else R_ItoL.copy()
        self.p_IinL = np.zeros(3) if p_IinL is None else p_IinL.copy()
        self.calib_idx = -1
        self.P = P0.copy()
        if calib_cov is not None:  # extrinsic calibration in the state (do_calib_ext)
            n = self.P.shape[0]
            P = np.zeros((n + 6, n + 6))
            P[:n, :n] = self.P
            P[n:, n:] = calib_cov
            self.P = P
            self.calib_idx = n
        self.intr_err_ori, self.intr_err_pos = intr_err_ori, intr_err_pos
        self._poly = {}

    # ---- bookkeeping (State.cpp:936-986) ------------------------------------------------------------------------
    def clone_times(self):
        return list(self.clones.keys())

    def oldest_clone_time(self):
        return next(iter(self.clones)) if self.clones else -np.inf

    def newest_clone_time(self):
        return next(reversed(self.clones)) if self.clones else -np.inf

    def oldest_3rd_clone_time(self):
        ts = self.clone_times()
        return ts[2] if len(ts) >= 3 else -np.inf

    def have_clone(self, t):
        return t in self.clones

    # ---- IMU propagation -------------------------------------------------------------------------------------------
    def propagate(self, wm, am, dt, noises):
        """one IMU interval of length dt with (already averaged) measurements wm, am. noises = (sg, sa, swg, swa)."""
        if dt <= 0:
            return
        sg, sa, swg, swa = noises
        w, a = wm - self.bg, am - self.ba
        R = self.R
        dR = exp_so3(-w * dt)
        a_G = R.T @ a - self.g
        F = np.eye(15)
        Jr = Jl_so3(w * dt)  # Jr(-w dt) = Jl(w dt)
        F[0:3, 0:3] = dR
        F[0:3, 9:12] = -dR @ Jr * dt
        F[3:6, 0:3] = -0.5 * R.T @ skew(a * dt * dt)
        F[3:6, 6:9] = np.eye(3) * dt
        F[3:6, 12:15] = -0.5 * R.T * dt * dt
        F[6:9, 0:3] = -R.T @ skew(a * dt)
        F[6:9, 12:15] = -R.T * dt
        G = np.zeros((15, 12))
        G[0:3, 0:3] = -dR @ Jr * dt
        G[3:6, 3:6] = -0.5 * R.T * dt * dt
        G[6:9, 3:6] = -R.T * dt
        G[9:12, 6:9] = np.eye(3) * dt
        G[12:15, 9:12] = np.eye(3) * dt
        Qc = np.diag(np.r_[np.full(3, sg * sg / dt), np.full(3, sa * sa / dt), np.full(3, swg * swg / dt), np.full(3, swa * swa / dt)])
        Qd = G @ Qc @ G.T
        # mean
        self.p = self.p + self.v * dt + 0.5 * a_G * dt * dt
        self.v = self.v + a_G * dt
        self.R = dR @ R
        # covariance: only the IMU rows/columns change
        P = self.P
        P[:15, :] = F @ P[:15, :]
        P[:, :15] = P[:, :15] @ F.T
        P[:15, :15] += Qd
        self.P = 0.5 * (P + P.T)
        self.time += dt

    # ---- stochastic cloning / marginalisation (StateHelper::augment_clone / marginalize) ------------------------------
    def augment_clone(self):
        n = self.P.shape[0]
        J = np.zeros((6, n))
        J[0:3, 0:3] = np.eye(3)
        J[3:6, 3:6] = np.eye(3)
        P = np.zeros((n + 6, n + 6))
        P[:n, :n] = self.P
        P[n:, :n] = J @ self.P
        P[:n, n:] = P[n:, :n].T
        P[n:, n:] = J @ self.P @ J.T
        self.P = P
        self.clones[self.time] = Clone(self.R, self.p, n)
        self._poly = {}

    def marginalize_oldest(self):
        t = self.oldest_clone_time()
        c = self.clones.pop(t)
        keep = np.r_[0:c.idx, c.idx + 6:self.P.shape[0]]
        self.P = self.P[np.ix_(keep, keep)]
        for o in self.clones.values():
            if o.idx > c.idx:
                o.idx -= 6
        self._poly = {}

    # ---- polynomial interpolation (State.cpp:428-520, 601-643, 675-826) ------------------------------------------------
    def _bounding(self, t):
        """intr_order+1 clone times around t (State::bounding_poses_n): the bracketing pair in the middle when possible."""
        ts = self.clone_times()
        n = self.intr_order + 1
        if len(ts) < n or t < ts[0] or t > ts[-1]:
            return None
        k = int(np.searchsorted(ts, t, side="right") - 1)  # ts[k] <= t
        k = min(k, len(ts) - 2)
        start = k - (n // 2 - 1)
        start = max(0, min(start, len(ts) - n))
        return ts[start:start + n]

    def _poly_data(self, times, fej):
        key = (times[0], fej)
        if key in self._poly:
            return self._poly[key]
        o = self.intr_order
        cl = [self.clones[t] for t in times]
        R0 = cl[0].R_fej if fej else cl[0].R
        p0 = cl[0].p_fej if fej else cl[0].p
        d_ori, d_pos = np.zeros(3 * o), np.zeros(3 * o)
        V = np.zeros((3 * o, 3 * o))
        JlInv, dbth = [], np.zeros((3 * o, 3))
        for i in range(1, o + 1):
            Ri = cl[i].R_fej if fej else cl[i].R
            pi = cl[i].p_fej if fej else cl[i].p
            Roi = Ri @ R0.T
            th = log_so3(Roi)
            d_ori[3 * (i - 1):3 * i] = th
            d_pos[3 * (i - 1):3 * i] = pi - p0
            dtp = times[i] - times[0]
            for j in range(1, o + 1):
                V[3 * (i - 1):3 * i, 3 * (j - 1):3 * j] = dtp**j * np.eye(3)
            Ji = np.linalg.inv(Jl_so3(th))
            JlInv.append(Ji)
            dbth[3 * (i - 1):3 * i] = Ji @ Roi
        Vinv = np.linalg.inv(V)
        data = dict(Vinv=Vinv, c_ori=Vinv @ d_ori, c_pos=Vinv @ d_pos, JlInv=JlInv, dbth=dbth,
                    dbp=np.tile(-np.eye(3), (o, 1)), R0=R0, p0=p0)
        self._poly[key] = data
        return data

    def get_interpolated_pose(self, t, fej=False):
        """State::get_interpolated_pose_poly (State.cpp:601-643) -> (R_GtoI, p_IinG) or None."""
        if t in self.clones:
            c = self.clones[t]
            return (c.R_fej, c.p_fej) if fej else (c.R, c.p)
        times = self._bounding(t)
        if times is None:
            return None
        d = self._poly_data(times, fej)
        o = self.intr_order
        dtm = t - times[0]
        A_ori = sum(dtm**(i + 1) * d["c_ori"][3 * i:3 * i + 3] for i in range(o))
        A_pos = sum(dtm**(i + 1) * d["c_pos"][3 * i:3 * i + 3] for i in range(o))
        return exp_so3(A_ori) @ d["R0"], d["p0"] + A_pos

    def get_interpolated_jacobian(self, t):
        """State::get_interpolated_jacobian (State.cpp:675-826) -> (R_fej, p_fej, [(clone_time, 6x6 dTdx)]) or None."""
        if t in self.clones:  # exactly at a clone (State::bounding_poses_n returns the clone itself)
            c = self.clones[t]
            return c.R_fej, c.p_fej, [(t, np.eye(6))]
        times = self._bounding(t)
        if times is None:
            return None
        d = self._poly_data(times, True)
        o = self.intr_order
        dtm = t - times[0]
        m_t = np.zeros((3, 3 * o))
        for i in range(o):
            m_t[:, 3 * i:3 * i + 3] = dtm**(i + 1) * np.eye(3)
        A_ori, A_pos = m_t @ d["c_ori"], m_t @ d["c_pos"]
        R_0toi = exp_so3(A_ori)
        R, p = R_0toi @ d["R0"], d["p0"] + A_pos
        mV = m_t @ d["Vinv"]
        dth_db = -Jl_so3(A_ori) @ mV
        out = []
        H0 = np.zeros((6, 6))
        H0[0:3, 0:3] = dth_db @ d["dbth"] + R_0toi
        H0[3:6, 3:6] = np.eye(3) + mV @ d["dbp"]
        out.append((times[0], H0))
        for i in range(o):
            Hi = np.zeros((6, 6))
            Hi[0:3, 0:3] = -dth_db[:, 3 * i:3 * i + 3] @ d["JlInv"][i]
            Hi[3:6, 3:6] = mV[:, 3 * i:3 * i + 3]
            out.append((times[i + 1], Hi))
        return R, p, out

    # ---- covariance access / update (StateHelper.cpp:227-254, 121-200) -------------------------------------------------
    def var_cols(self, var):
        """var: ('clone', t) or ('calib',) -> column indices"""
        if var[0] == "clone":
            i = self.clones[var[1]].idx
            return np.arange(i, i + 6)
        return np.arange(self.calib_idx, self.calib_idx + 6)

    def get_marginal_covariance(self, H_order):
        idx = np.concatenate([self.var_cols(v) for v in H_order])
        return self.P[np.ix_(idx, idx)]

    def ekf_update(self, H_order, H, res):
        """StateHelper::EKFUpdate with R = I. Returns False (and changes nothing) if a diagonal would go negative."""
        idx = np.concatenate([self.var_cols(v) for v in H_order])
        M = self.P[:, idx] @ H.T
        S = H @ self.P[np.ix_(idx, idx)] @ H.T + np.eye(len(res))
        K = np.linalg.solve(S, M.T).T
        Pn = self.P - K @ M.T
        if (np.diag(Pn) < 0).any():
            return False
        self.P = 0.5 * (Pn + Pn.T)
        self.apply(K @ res)
        return True

    def apply(self, dx):
        self.R = exp_so3(-dx[0:3]) @ self.R
        self.p = self.p + dx[3:6]
        self.v = self.v + dx[6:9]
        self.bg = self.bg + dx[9:12]
        self.ba = self.ba + dx[12:15]
        if self.calib_idx >= 0:
            i = self.calib_idx
            self.R_ItoL = exp_so3(-dx[i:i + 3]) @ self.R_ItoL
            self.p_IinL = self.p_IinL + dx[i + 3:i + 6]
        for c in self.clones.values():
            c.R = exp_so3(-dx[c.idx:c.idx + 3]) @ c.R
            c.p = c.p + dx[c.idx + 3:c.idx + 6]
        self._poly = {k: v for k, v in self._poly.items() if k[1]}  # fej polynomials stay valid
