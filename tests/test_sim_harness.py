"""Closed-loop harness, CPU part: the restated simulator pieces and EKF core are self-consistent, and an
IMU+LiDAR run with the CPU-oracle LiDAR updater tracks the ground truth while IMU-only drifts."""
import numpy as np
import pytest

from mins_b200.sim import se3
from mins_b200.sim.ekf import State
from mins_b200.sim.traj import BsplineSE3, synthetic_control_poses
from mins_b200.sim.world import PlaneWorld, lidar_ray_dirs, simulate_scan, synthetic_floorplan


def test_so3_se3_helpers():
    rng = np.random.default_rng(0)
    for _ in range(50):
        w = rng.normal(size=3)
        w = w / np.linalg.norm(w) * rng.choice([1e-9, 0.3, 2.5])  # principal range |w| < pi
        R = se3.exp_so3(w)
        assert np.abs(R @ R.T - np.eye(3)).max() < 1e-12 and np.abs(se3.log_so3(R) - w).max() < 1e-9
        q = se3.rot_to_quat_jpl(R)
        assert np.abs(se3.quat_jpl_to_rot(q) - R).max() < 1e-12
        xi = np.r_[w, rng.normal(size=3)]
        assert np.abs(se3.log_se3(se3.exp_se3(xi)) - xi).max() < 1e-9
        # Jl is the derivative of exp: exp(w + d) ~ exp(Jl d) exp(w)
        d = 1e-6 * rng.normal(size=3)
        assert np.abs(se3.exp_so3(w + d) - se3.exp_so3(se3.Jl_so3(w) @ d) @ R).max() < 1e-10


def test_bspline_is_smooth_and_kinematics_are_consistent():
    sp = BsplineSE3(synthetic_control_poses(0, 20.0))
    t = sp.t_start + 3.21
    R, p, w, v, a = sp.get_kinematics(t)
    h = 1e-4
    _, p1 = sp.get_pose(t + h)
    _, p0 = sp.get_pose(t - h)
    assert np.abs((p1 - p0) / (2 * h) - v).max() < 1e-5
    R1, _ = sp.get_pose(t + h)
    assert np.abs(se3.exp_so3(-w * h) @ R - R1).max() < 1e-6          # JPL kinematics R(t+h) = exp(-w h) R(t)
    # C2 continuity across a control-point boundary
    tb = sp.ctrl_t[10]
    a_l, a_r = sp.get_kinematics(tb - 3e-3)[4], sp.get_kinematics(tb + 3e-3)[4]
    assert np.abs(a_l - a_r).max() < 0.05 * max(1.0, np.abs(a_l).max())
    assert 0.5 < np.linalg.norm(v) < 5.0


def test_raycast_against_closed_form_box():
    """a 10 x 6 room, z in [-1, 4]: rays from the centre hit walls/floor/ceiling at analytically known ranges"""
    w = PlaneWorld(np.array([[-5, -3, -5, 3], [-5, -3, 5, -3], [-5, 3, 5, 3], [5, -3, 5, 3.0]]))
    o = np.array([0.3, -0.2, 0.5])
    d = lidar_ray_dirs([-60, -20, 0, 15, 70], -180, 180, 7.0)
    t = w.raycast(o, d)
    lo, hi = np.array([-5, -3, -1.0]), np.array([5, 3, 4.0])
    with np.errstate(divide="ignore"):
        tt = np.where(d > 0, (hi - o) / d, np.where(d < 0, (lo - o) / d, np.inf))
    assert np.allclose(t, tt.min(1), rtol=1e-12)
    pts = simulate_scan(w, np.eye(3), o, np.eye(3), np.zeros(3), d, 1.0, 100.0, 0.0, np.random.default_rng(0))
    assert pts.dtype == np.float32 and len(pts) == (tt.min(1) >= 1.0).sum()
    assert np.allclose(np.linalg.norm(pts, axis=1), tt.min(1)[tt.min(1) >= 1.0], rtol=1e-6)


def make_state(rng, n_clones=8, dt=0.05):
    sp = BsplineSE3(synthetic_control_poses(3, 10.0))
    st = State(np.eye(3), np.zeros(3), np.zeros(3), np.zeros(3), np.zeros(3), np.eye(15) * 1e-4)
    for k in range(n_clones):
        t = round(sp.t_start + 1.0 + k * dt, 6)
        st.R, st.p = sp.get_pose(t)
        st.time = t
        st.augment_clone()
    return st


@pytest.mark.parametrize("frac", [0.0, 0.3, 0.77, 2.4, 5.6])
def test_interpolation_jacobian_matches_finite_differences(frac):
    """State::get_interpolated_jacobian (State.cpp:675-826) restated: d(pose at t)/d(clone poses), JPL error state"""
    rng = np.random.default_rng(1)
    st = make_state(rng)
    ts = st.clone_times()
    t = ts[1] + frac * 0.05
    R, p, blocks = st.get_interpolated_jacobian(t)
    Re, pe = st.get_interpolated_pose(t)
    assert np.abs(R - Re).max() < 1e-12 and np.abs(p - pe).max() < 1e-12   # fej == est right after cloning
    assert len(blocks) == (1 if frac == 0.0 else 4)
    eps = 1e-6
    for tc, D in blocks:
        for j in range(6):
            c = st.clones[tc]
            R_keep, p_keep = c.R.copy(), c.p.copy()
            d = np.zeros(6)
            d[j] = eps
            c.R = se3.exp_so3(-d[:3]) @ c.R                                  # R_true = exp(-dth) R_hat
            c.p = c.p + d[3:]
            st._poly = {}
            Rn, pn = st.get_interpolated_pose(t)
            c.R, c.p = R_keep, p_keep
            st._poly = {}
            num = np.r_[-se3.log_so3(Rn @ Re.T), pn - pe] / eps
            assert np.abs(num - D[:, j]).max() < 2e-5, (tc, j, num, D[:, j])
    # outside the clone window: no interpolation (the scan waits, UpdaterLidar.cpp:190-191)
    assert st.get_interpolated_jacobian(ts[0] - 0.01) is None and st.get_interpolated_pose(ts[-1] + 0.01) is None


def test_clone_marginalisation_keeps_covariance_consistent():
    rng = np.random.default_rng(2)
    st = make_state(rng, n_clones=5)
    n = st.P.shape[0]
    assert n == 15 + 30 and [c.idx for c in st.clones.values()] == [15, 21, 27, 33, 39]
    assert np.abs(st.P[15:21, 15:21] - st.P[0:6, 0:6]).max() < 1e-15           # a clone is a copy of the IMU pose
    st.marginalize_oldest()
    assert st.P.shape[0] == n - 6 and [c.idx for c in st.clones.values()] == [15, 21, 27, 33]
    H = rng.normal(size=(3, 12))
    order = [("clone", st.clone_times()[0]), ("clone", st.clone_times()[2])]
    P_before = st.P.copy()
    assert st.ekf_update(order, H, rng.normal(size=3) * 1e-3)
    assert (np.linalg.eigvalsh(P_before - st.P) > -1e-12).all()                # information only increases


def test_closed_loop_lidar_with_cpu_oracle_tracks_ground_truth():
    from mins_b200.sim.runner import SimConfig, run
    from oracle.backend import OracleBackend
    r_imu = run(SimConfig(seed=4, duration=8.0, use_lidar=False), OracleBackend())
    r_ldr = run(SimConfig(seed=4, duration=8.0, use_lidar=True), OracleBackend(threads=4))
    assert r_ldr["n_updates"] >= 70 and r_ldr["n_acc"] > 0.5 * r_ldr["n_points"]
    assert r_ldr["rmse_pos_m"] < 0.05 and r_ldr["rmse_ori_deg"] < 0.6
    assert r_ldr["final_pos_err"] < 0.5 * r_imu["final_pos_err"]
    assert 0.05 < r_ldr["nees_pos"] < 5.0
