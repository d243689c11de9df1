"""CPU checks of the simulator-LiDAR oracle (oracle/pyoracle.py::sim_lidar_scan, philox4x32): the Philox-4x32-10 known-answer
vector of the Random123 distribution, and agreement between the oracle's literal restatement of
SimulationPlane::calculate_intersection and the host harness' own ray-caster (mins_b200/sim/world.py) — two independent
restatements of mins/src/sim/Simulator.cpp:742-815."""
import numpy as np


def test_philox_known_answer(po):
    # Random123 kat_vectors: philox4x32-10, counter 0 0 0 0, key 0 0
    w = po.philox4x32(np.array([0]), 0, 0)[0]
    assert [int(x) for x in w] == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    n = po.sim_noise(np.arange(100000), 99)
    assert abs(n.mean()) < 0.01 and abs(n.std() - 1) < 0.01


def test_oracle_raycast_matches_harness_raycast(po):
    from mins_b200.sim.se3 import exp_so3
    from mins_b200.sim.world import PlaneWorld, lidar_ray_dirs, simulate_scan, synthetic_floorplan
    world = PlaneWorld(synthetic_floorplan(0))
    dirs = lidar_ray_dirs(np.arange(-80, 81, 10), -180, 180, 2.0)
    R_GtoI = exp_so3(np.array([0.05, -0.1, 0.7]))
    p_IinG = np.array([1.0, -2.0, 0.8])
    R_ItoL, p_IinL = exp_so3(np.array([0.26, 0, 0])), np.array([-0.1, -0.12, -0.07])
    pts = simulate_scan(world, R_GtoI, p_IinG, R_ItoL, p_IinL, dirs, 0.1, 100.0, 0.0, None)
    p_LinG = p_IinG + R_GtoI.T @ (-R_ItoL.T @ p_IinL)
    want, keep, best = po.sim_lidar_scan(world.quads, p_LinG, R_GtoI.T @ R_ItoL.T, dirs, 0.1, 100.0)
    assert len(pts) == len(want) == len(keep) > 2000
    assert np.abs(pts - want[:, :3]).max() < 1e-6
    # every kept ray hits inside the hall (floor plan 40 m x 25 m, z in [-1, 4])
    assert best[keep].max() < 50.0
