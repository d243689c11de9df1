"""The closed-loop harness in host C++ (mins_b200/host/sim/*): the minimal EKF core + mins_gpu::UpdaterLidar.
CPU part: the IMU + clone-window core alone (no device) against the Python harness it replaces, noise free, on the MINS
simulation's own UD_Warehouse trajectory. GPU part: the whole loop through the C++ shim and the C ABI."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

BIN = os.path.join(ROOT, "mins_b200", "host", "mins_sim_run")
BSPLINE = os.path.join(ROOT, "tests", "golden", "UD_Warehouse_bspline.txt")
FLOOR = os.path.join(ROOT, "tests", "golden", "UD_Warehouse_floorplan.txt")


def run_cpp(*args, traj=None):
    cmd = [BIN, "--bspline", BSPLINE, "--floorplan", FLOOR, *map(str, args)]
    if traj:
        cmd += ["--traj", traj]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    return json.loads(p.stdout.strip().splitlines()[-1])


def test_fixtures_are_the_reference_files():
    """253 control poses over 52 s; 188 wall segments in a 77.4 m x 54.1 m hall at 0.1 m per file unit (Simulator.cpp:990-1096)"""
    from mins_b200.sim.traj import BsplineSE3, load_bspline_dataset
    from mins_b200.sim.world import PlaneWorld, load_floorplan
    tp = load_bspline_dataset(BSPLINE)
    assert tp.shape == (253, 8) and tp[0, 0] == 0.0 and tp[-1, 0] == 52.0
    seg = load_floorplan(FLOOR)
    assert seg.shape == (188, 4) and seg.max() == pytest.approx(77.4) and seg[:, [1, 3]].max() == pytest.approx(54.1)
    assert len(PlaneWorld(seg).quads) == 190
    sp = BsplineSE3(tp)
    assert sp.t_end - sp.t_start > 50.0


def test_cpp_core_matches_the_python_harness_imu_only(tmp_path):
    from mins_b200.sim.runner import SimConfig, run
    traj = str(tmp_path / "traj.txt")
    r = run_cpp("--no-lidar", "--no-noise", "--duration", 6, traj=traj)
    cfg = SimConfig(seed=0, duration=6.0, use_lidar=False, remove_noise=True, bspline_path=BSPLINE, floorplan_path=FLOOR)
    py = run(cfg, None)
    c = np.loadtxt(traj)
    assert c.shape == py["traj"].shape and r["n_clones"] == len(c) == 120
    assert np.abs(c[:, 0] - py["traj"][:, 0]).max() < 1e-9
    assert np.abs(c[:, 1:4] - py["traj"][:, 1:4]).max() < 1e-7          # positions [m]
    assert np.abs(c[:, 4:] - py["traj"][:, 4:]).max() < 1e-9            # JPL quaternions
    assert r["rmse_pos_m"] == pytest.approx(py["rmse_pos_m"], rel=1e-6) and r["rmse_ori_deg"] == pytest.approx(py["rmse_ori_deg"], rel=1e-6)
    assert r["nees_pos"] == pytest.approx(py["nees_pos"], rel=1e-6)     # same covariance propagation and cloning


def test_cpp_harness_writes_the_state_logger_trajectory_format(tmp_path):
    """`--state-log`: one line per clone in the format of State_Logger::save_trajectory_to_file (mins/src/utils/State_Logger.h:323-340),
    which mins_eval's trajectory tools load: time, position, JPL quaternion (6 significant digits), the upper triangles of the
    orientation and the position blocks of the IMU pose covariance (10 significant digits) - 20 columns."""
    traj, log = str(tmp_path / "traj.txt"), str(tmp_path / "state.txt")
    run_cpp("--no-lidar", "--duration", 3, "--seed", 2, "--state-log", log, traj=traj)
    t, s = np.loadtxt(traj), np.loadtxt(log)
    assert s.shape == (len(t), 20) and len(t) == 60
    assert np.abs(s[:, :8] - t).max() < 5e-6 * np.abs(t).max()                      # the same poses at the reference's precision
    ori, pos = s[:, 8:14], s[:, 14:20]
    for blk in (ori, pos):                                                          # [00 01 02 11 12 22] of a covariance block
        C = np.stack([blk[:, [0, 1, 2]], blk[:, [1, 3, 4]], blk[:, [2, 4, 5]]], 1)
        assert (np.linalg.eigvalsh(C) > 0).all()
    assert pos[-1, 0] > pos[0, 0] and pos[-1, 5] > pos[0, 5]                        # IMU only: the position uncertainty grows


def test_monte_carlo_launcher_shards_seeds_over_persistent_workers():
    """scripts/mc_sim_cpp.py (BASELINE config "64 independent runs sharded over 1/2/4/8 B200"): one worker process per (GPU, slot),
    seed s on worker s mod workers, every worker runs its seeds in one process. IMU-only here (no device): 7 runs over 2 x 2 workers."""
    import sys
    p = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "mc_sim_cpp.py"), "--gpus", "2", "--per-gpu", "2", "--runs", "7",
                        "--duration", "1", "--seed0", "10", "--no-lidar"], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    summary = json.loads(p.stdout.strip().splitlines()[-1])
    runs = [json.loads(line) for line in p.stderr.splitlines() if line.startswith("{")]
    assert summary["runs"] == 7 and summary["gpus"] == 2 and summary["per_gpu"] == 2
    assert sorted(r["seed"] for r in runs) == list(range(10, 17))                    # every seed exactly once
    by_gpu = {g: sorted(r["seed"] for r in runs if r["gpu"] == g) for g in (0, 1)}
    assert by_gpu[0] == [10, 12, 14, 16] and by_gpu[1] == [11, 13, 15]                # workers 0..3 = (gpu0,s0) (gpu1,s0) (gpu0,s1) (gpu1,s1)
    assert all(r["n_scans"] == 0 and r["n_clones"] == 20 for r in runs)


def test_cpp_harness_needs_a_device_for_lidar():
    """no CPU fallback: without a CUDA device the LiDAR run fails loudly"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU box")
    p = subprocess.run([BIN, "--bspline", BSPLINE, "--floorplan", FLOOR, "--duration", "1"], capture_output=True, text=True, timeout=120)
    assert p.returncode != 0 and ("ml_create" in p.stderr or "CUDA" in p.stderr)


@pytest.mark.gpu
def test_cpp_closed_loop_on_ud_warehouse_matches_the_python_mirror(tmp_path):
    """the same noise-free run through the C++ shim (mins_gpu::UpdaterLidar) and through the Python mirror over the same C ABI:
    same scans (device ray-caster), same updates, same trajectory"""
    from mins_b200.lidar import OptionsLidar
    from mins_b200.sim.runner import SimConfig, run
    from mins_b200.sim.updater import GpuBackend
    traj = str(tmp_path / "traj.txt")
    r = run_cpp("--no-noise", "--duration", 8, traj=traj)
    gpu = GpuBackend(max_scan_points=1 << 17, max_map_points=1 << 19)
    try:
        py = run(SimConfig(seed=0, duration=8.0, remove_noise=True, scan_source="cuda", use_gt_lidar=True, bspline_path=BSPLINE,
                           floorplan_path=FLOOR, op=OptionsLidar(raw_do_downsample=True, map_use_icp=False)), gpu)
    finally:
        gpu.close()
    c = np.loadtxt(traj)
    assert r["n_updates"] == py["n_updates"] >= 70 and r["n_scans"] == py["n_scans"]
    assert np.abs(c[:, 1:4] - py["traj"][:, 1:4]).max() < 1e-5
    assert r["rmse_pos_m"] < 0.01 and py["rmse_pos_m"] < 0.01


@pytest.mark.gpu
def test_cpp_closed_loop_with_noise_is_accurate():
    """IMU + LiDAR on UD_Warehouse with the simulation YAML's noises: ATE in the band of the reference's published IMU+LiDAR
    numbers (mins_eval/ReadMe.md:153-154: 0.24-0.32 deg / 0.037-0.041 m on its example datasets)"""
    rs = [run_cpp("--seed", s, "--duration", 20) for s in (0, 1, 2)]
    for r in rs:
        assert r["n_updates"] >= 0.9 * r["n_scans"] and r["n_acc"] > 0
        assert r["rmse_pos_m"] < 0.08 and r["rmse_ori_deg"] < 0.6
        assert r["ldr_wall_s"] < r["wall_s"]
