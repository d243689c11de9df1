"""GPU parity of the simulator LiDAR model: ml_sim_raycast (rays x planes nearest hit, range gate, fp32
points, counter-based noise, stable compaction) against the numpy restatement of Simulator::get_lidar_pointcloud /
SimulationPlane::calculate_intersection in oracle/pyoracle.py.

Bars: identical kept-ray set and order; points within 2e-6 m noise-free (fp64 hit distance, evaluation order differs by FMA
contraction, then one fp32 rounding) and within 3e-6 m with noise (logf/cosf/sinf are not correctly rounded on either side).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pose(seed):
    from mins_b200.sim.se3 import exp_so3
    rng = np.random.default_rng(seed)
    R = exp_so3(rng.normal(0, 0.3, 3))
    p = np.array([rng.uniform(-8, 8), rng.uniform(-4, 4), rng.uniform(0.2, 2.0)])
    return p, R


@pytest.fixture()
def sim_map(cuda_ok):
    from mins_b200.lidar import LidarMap
    m = LidarMap(max_scan_points=1 << 17, max_map_points=1 << 18)
    yield m
    m.close()


@pytest.mark.parametrize("pattern", ["sim17x181", "beam64x2048"])
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_raycast_matches_oracle_noise_free(sim_map, po, pattern, seed):
    from mins_b200.sim.world import PlaneWorld, lidar_ray_dirs, synthetic_floorplan
    world = PlaneWorld(synthetic_floorplan(seed))
    if pattern == "sim17x181":
        dirs = lidar_ray_dirs(np.arange(-80, 81, 10), -180, 180, 2.0)   # config_lidar.yaml:38-44
    else:
        dirs = lidar_ray_dirs(np.linspace(-24.8, 2.0, 64), -180, 180 - 360 / 2048, 360 / 2048)
    p, R = _pose(seed)
    sim_map.sim_set_world(world.packed())
    sim_map.sim_set_rays(dirs)
    got = sim_map.sim_raycast(p, R, 0.5, 30.0)
    want, keep, best = po.sim_lidar_scan(world.quads, p, R, dirs, 0.5, 30.0)
    assert len(keep) > len(dirs) // 3
    assert len(got) == len(want)
    assert np.abs(got - want).max() <= 2e-6
    assert np.all(got[:, 3] == 0)


def test_raycast_noise_matches_oracle_and_is_deterministic(sim_map, po):
    from mins_b200.sim.world import PlaneWorld, lidar_ray_dirs, synthetic_floorplan
    world = PlaneWorld(synthetic_floorplan(3))
    dirs = lidar_ray_dirs(np.arange(-80, 81, 10), -180, 180, 1.0)
    p, R = _pose(7)
    sim_map.sim_set_world(world.packed())
    sim_map.sim_set_rays(dirs)
    a = sim_map.sim_raycast(p, R, 0.1, 100.0, noise=0.01, seed=(5 << 32) | 77)
    b = sim_map.sim_raycast(p, R, 0.1, 100.0, noise=0.01, seed=(5 << 32) | 77)
    c = sim_map.sim_raycast(p, R, 0.1, 100.0, noise=0.01, seed=78)
    clean = sim_map.sim_raycast(p, R, 0.1, 100.0)
    want, keep, _ = po.sim_lidar_scan(world.quads, p, R, dirs, 0.1, 100.0, noise=0.01, seed=(5 << 32) | 77)
    assert np.array_equal(a, b) and len(a) == len(c) == len(clean) == len(want)
    assert not np.array_equal(a, c)
    assert np.abs(a - want).max() <= 3e-6
    z = (a[:, :3] - clean[:, :3]) / np.float32(0.01)       # the injected deviates
    assert abs(z.mean()) < 0.02 and abs(z.std() - 1.0) < 0.02 and np.abs(z).max() < 6.0
    assert abs(np.corrcoef(z[:, 0], z[:, 1])[0, 1]) < 0.03


def test_raycast_scan_feeds_the_update_in_place(sim_map, po, workload):
    """the ray-cast scan stays in HBM: ml_scan_update on the device pointer == ml_scan_update on the downloaded copy."""
    from mins_b200 import capi
    from mins_b200.sim.world import PlaneWorld, lidar_ray_dirs, synthetic_floorplan
    w = workload("G-small")
    op = w["op"]
    m = sim_map
    m.Build(w["map"])
    world = PlaneWorld(synthetic_floorplan(0))
    dirs = lidar_ray_dirs(np.arange(-80, 81, 10), -180, 180, 2.0)
    m.sim_set_world(world.packed())
    m.sim_set_rays(dirs)
    host = m.sim_raycast(np.array([0.0, 0.0, 1.0]), np.eye(3), op.min_range, op.max_range, noise=0.01, seed=1)
    ptr, n = m.sim_scan_device()
    assert n == len(host) > 1000
    hx = w["total_hx"]
    T = np.ascontiguousarray(w["T_LtoM"], np.float32)
    P = np.ascontiguousarray(w["P"], np.float64)

    def upd(xyz_ptr):
        HtH, Htr, cnt = np.zeros((hx, hx)), np.zeros(hx), capi.ml_counters()
        rc = m.lib.ml_scan_update(m.h, xyz_ptr, 4, n, w["t_rel"], op.min_range, op.max_range, 0.0, T.ctypes.data,
                                  C.byref(w["consts"]), P.ctypes.data, HtH.ctypes.data, Htr.ctypes.data, C.byref(cnt))
        assert rc in (0, 1)
        return HtH, Htr, cnt.as_dict()

    dev = upd(ptr)
    kept_dev = m.scan(2)
    hst = upd(host.ctypes.data)
    assert np.array_equal(kept_dev, m.scan(2))
    assert dev[2] == hst[2] and np.array_equal(dev[0], hst[0]) and np.array_equal(dev[1], hst[1])


def test_raycast_with_more_planes_than_one_shared_memory_pass(sim_map, po):
    """> 256 planes: k_ray_cast stages the world in several shared-memory passes; plus the empty-result edge case."""
    from mins_b200.sim.world import PlaneWorld, lidar_ray_dirs
    rng = np.random.default_rng(5)
    segs = []
    for _ in range(90):  # 90 boxes x 4 walls = 360 wall planes (+ floor and ceiling)
        cx, cy, s = rng.uniform(-30, 30), rng.uniform(-30, 30), rng.uniform(0.3, 1.5)
        if abs(cx) < 3 and abs(cy) < 3:
            cx += 6
        segs += [[cx - s, cy - s, cx - s, cy + s], [cx - s, cy - s, cx + s, cy - s], [cx - s, cy + s, cx + s, cy + s], [cx + s, cy - s, cx + s, cy + s]]
    world = PlaneWorld(np.array(segs))
    assert len(world.quads) > 256
    dirs = lidar_ray_dirs(np.arange(-30, 31, 5), -180, 180, 0.5)
    p, R = np.array([0.0, 0.0, 0.5]), np.eye(3)
    sim_map.sim_set_world(world.packed())
    sim_map.sim_set_rays(dirs)
    got = sim_map.sim_raycast(p, R, 0.5, 100.0)
    want, keep, best = po.sim_lidar_scan(world.quads, p, R, dirs, 0.5, 100.0)
    assert len(got) == len(want) > 1000 and np.abs(got - want).max() <= 4e-6
    # a range gate nothing passes
    none = sim_map.sim_raycast(p, R, 1000.0, 2000.0)
    assert none.shape == (0, 4)


def test_motion_distorted_scan_is_undone_by_deskew(sim_map):
    """§8f N4 with the simulator's distortion model (Simulator.cpp:761-771): a scan ray-cast from a moving sensor
    (ml_sim_raycast_sweep) and de-skewed with the same pose track (ml_scan_deskew, the loop body of LidarHelper.cpp:71-85)
    lands on the world points expressed in the end-of-sweep frame."""
    from mins_b200.lidar import OptionsLidar
    from mins_b200.sim.se3 import exp_so3
    from mins_b200.sim.world import PlaneWorld, lidar_ray_dirs, synthetic_floorplan
    world = PlaneWorld(synthetic_floorplan(0))
    dirs = lidar_ray_dirs(np.arange(-80, 81, 10), -180, 180, 2.0)
    n, K = len(dirs), 33
    R0, p0 = exp_so3(np.array([0.02, -0.01, 0.3])), np.array([0.5, -1.0, 1.0])
    w, v = np.array([0.05, -0.1, 0.6]), np.array([1.0, 0.4, 0.05])           # rad and m over the sweep
    Rk = [R0 @ exp_so3(w * s) for s in np.linspace(0, 1, K)]
    pk = [p0 + v * s for s in np.linspace(0, 1, K)]
    knots = np.array([np.r_[R.ravel(), p] for R, p in zip(Rk, pk)])
    sim_map.sim_set_world(world.packed())
    sim_map.sim_set_rays(dirs)
    skewed = sim_map.sim_raycast_sweep(knots, 0.01, 200.0)
    static = sim_map.sim_raycast(pk[-1], Rk[-1], 0.01, 200.0)
    assert len(skewed) == len(static) == n                                    # closed hall: every ray returns
    # ground truth on the host: the world point of ray i, expressed in the end frame
    u = np.arange(n) / n * (K - 1)
    k0 = np.minimum(u.astype(int), K - 2)
    a = (u - k0)[:, None]
    P = (1 - a) * knots[k0] + a * knots[k0 + 1]
    Ri, oi = P[:, :9].reshape(n, 3, 3), P[:, 9:]
    world_pts = np.einsum("nij,nj->ni", Ri, skewed[:, :3].astype(np.float64)) + oi
    want = (world_pts - pk[-1]) @ Rk[-1]
    # estimator side: T_LitoL = [R_end^T R_i | R_end^T (p_i - p_end)] (LidarHelper.cpp:73-79), fp32 knots
    T = np.zeros((K, 3, 4), np.float32)
    for j in range(K):
        T[j, :, :3] = Rk[-1].T @ Rk[j]
        T[j, :, 3] = Rk[-1].T @ (pk[j] - pk[-1])
    op = OptionsLidar(min_range=0.01, max_range=200.0)
    assert sim_map.feed_scan(skewed, 0.0, op) == n
    sim_map.deskew(T.reshape(K, 12), n)
    got = sim_map.scan(0)[:, :3]
    assert np.abs(got - want).max() < 2e-4                                    # fp32 knots and transform at ~20 m range
    assert np.abs(skewed[:, :3] - want).max() > 0.3                           # without de-skewing the error is decimetres
    # every de-skewed point lies on a wall/floor/ceiling plane of the world
    pw = got.astype(np.float64) @ Rk[-1].T + pk[-1]
    A = world.abcd / np.linalg.norm(world.abcd[:, :3], axis=1, keepdims=True)
    assert np.abs(pw @ A[:, :3].T + A[:, 3]).min(1).max() < 3e-4
