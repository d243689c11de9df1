"""Closed-loop parity (BASELINE.json config 0/3): the SAME IMU+LiDAR estimator run twice on the same seeded
simulation, once with the GPU LiDAR updater (C ABI) and once with the CPU-oracle updater; the trajectories and their ATE
must agree. `map_use_icp: false` (the only deterministic configuration of the reference)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,downsample", [(1, True), (2, False)])
def test_gpu_and_oracle_updaters_give_the_same_trajectory(cuda_ok, seed, downsample):
    from mins_b200.lidar import OptionsLidar
    from mins_b200.sim.runner import SimConfig, run
    from mins_b200.sim.updater import GpuBackend
    from oracle.backend import OracleBackend
    op = OptionsLidar(raw_do_downsample=downsample, map_use_icp=False)
    cfg = SimConfig(seed=seed, duration=6.0, op=op)
    gpu = GpuBackend(max_scan_points=1 << 14, max_map_points=1 << 18)
    try:
        r_gpu = run(cfg, gpu)
    finally:
        gpu.close()
    r_cpu = run(cfg, OracleBackend())
    assert r_gpu["n_updates"] == r_cpu["n_updates"] >= 50
    assert r_gpu["n_points"] == r_cpu["n_points"]
    # identical accept/reject decisions except points within rounding of a gate (none expected; allow a handful)
    assert abs(r_gpu["n_acc"] - r_cpu["n_acc"]) <= 3
    d = np.abs(r_gpu["traj"][:, 1:4] - r_cpu["traj"][:, 1:4]).max()
    assert d < 1e-5, d                                                   # metres, over the whole run
    assert r_gpu["rmse_pos_m"] == pytest.approx(r_cpu["rmse_pos_m"], abs=1e-5)
    assert r_gpu["rmse_ori_deg"] == pytest.approx(r_cpu["rmse_ori_deg"], abs=1e-4)
    assert r_gpu["rmse_pos_m"] < 0.05 and r_gpu["map_points"] == r_cpu["map_points"]


@pytest.mark.parametrize("seed", [0, 5])
def test_gpu_and_oracle_agree_on_the_mins_simulation_itself(cuda_ok, seed):
    """BASELINE.json config[0] on the reference's own inputs: UD_Warehouse B-spline + floor plan (mins_data, committed under
    tests/golden), map seeded from the dense 180 x 361 ground-truth scan (Initializer.cpp:308-340), the simulation YAML's option
    set (mins/config/simulation/config_lidar.yaml, `map_use_icp: false` - the deterministic configuration). Same seeded scans
    through the GPU updater and through the CPU oracle updater in the same EKF: the trajectories must coincide."""
    from mins_b200.lidar import OptionsLidar
    from mins_b200.sim.runner import UD_WAREHOUSE, SimConfig, run
    from mins_b200.sim.updater import GpuBackend
    from oracle.backend import OracleBackend
    op = OptionsLidar(raw_do_downsample=True, map_use_icp=False)
    cfg = SimConfig(seed=seed, duration=5.0, op=op, bspline_path=UD_WAREHOUSE[0], floorplan_path=UD_WAREHOUSE[1], use_gt_lidar=True)
    gpu = GpuBackend(max_scan_points=1 << 17, max_map_points=1 << 19)
    try:
        r_gpu = run(cfg, gpu)
    finally:
        gpu.close()
    r_cpu = run(cfg, OracleBackend(threads=16))
    assert r_gpu["n_updates"] == r_cpu["n_updates"] >= 45 and r_gpu["n_points"] == r_cpu["n_points"]
    assert abs(r_gpu["n_acc"] - r_cpu["n_acc"]) <= 3
    d = np.abs(r_gpu["traj"][:, 1:4] - r_cpu["traj"][:, 1:4]).max()
    assert d < 1e-5, d
    assert r_gpu["rmse_pos_m"] == pytest.approx(r_cpu["rmse_pos_m"], abs=1e-5) and r_gpu["rmse_pos_m"] < 0.05
    assert r_gpu["map_points"] == r_cpu["map_points"] > 10000


def test_closed_loop_with_gpu_icp_and_gpu_raycaster(cuda_ok):
    """BASELINE.json config 0 with the simulation's own default `map_use_icp: true` (config_lidar.yaml:27) and the scans
    ray-cast on the device. The reference's ICP is non-deterministic, so GPU and oracle cannot follow the same trajectory;
    the bar is statistical: ICP converges on (nearly) every scan on both sides, in a similar number of iterations, and the
    two closed-loop runs end up equally accurate. (Measured: on this 17-ring LiDAR in the synthetic hall the restated
    libpointmatcher chain makes the run LESS accurate than the plain path, ~0.10 m vs ~0.04 m, and the GPU chain shows the
    same behaviour - ring-collinear neighbourhoods give poor normals; that is a property of the method, kept.)"""
    from mins_b200.lidar import OptionsLidar
    from mins_b200.sim.runner import SimConfig, run
    from mins_b200.sim.updater import GpuBackend
    from oracle.backend import OracleBackend
    op = OptionsLidar(raw_do_downsample=True, map_use_icp=True, map_icp_dist=50.0)
    gpu = GpuBackend(max_scan_points=1 << 14, max_map_points=1 << 18)
    try:
        r = run(SimConfig(seed=3, duration=4.0, op=op, scan_source="cuda"), gpu)
    finally:
        gpu.close()
    r_cpu = run(SimConfig(seed=3, duration=4.0, op=op), OracleBackend())
    for x in (r, r_cpu):
        assert x["icp_runs"] >= 35 and x["icp_ok"] >= 0.8 * x["icp_runs"]  # retries of failed alignments count as runs
        assert 4 < x["icp_iters"] / x["icp_runs"] < 25
        assert x["rmse_pos_m"] < 0.3 and x["n_updates"] >= 35
    assert 0.4 < r["rmse_pos_m"] / r_cpu["rmse_pos_m"] < 2.5
    assert 0.5 < (r["icp_iters"] / r["icp_runs"]) / (r_cpu["icp_iters"] / r_cpu["icp_runs"]) < 2.0
