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
