"""CPU checks of the ICP oracle (oracle/icp_oracle.py): the restated libpointmatcher chain recovers a known transform on the
simulated hall, its filters have the reference's documented statistics, and the chain reports failure like
LidarHelper.cpp:177-183 when it cannot converge."""
import numpy as np

from icp_scene import make_scene, pose_error


def test_filters_follow_the_reference_rules():
    from oracle import icp_oracle as io
    rng = np.random.default_rng(0)
    keep = io.random_sampling(10000, 0.75, rng)
    assert abs(len(keep) / 10000 - 0.75) < 0.02 and len(keep) <= int(np.float32(10000) * np.float32(0.75)) + 1
    pts = rng.normal(0, 10, (5000, 3)).astype(np.float32)
    assert np.array_equal(io.max_dist_filter(pts, 12.0), np.where(np.linalg.norm(pts, axis=1).astype(np.float32) < 12.0)[0])
    # a noisy plane z = 0.2 x: every bin normal is +-(−0.2, 0, 1)/|.| up to noise, about half of the points survive
    xy = rng.uniform(-5, 5, (4000, 2))
    plane = np.c_[xy, 0.2 * xy[:, 0] + rng.normal(0, 0.002, 4000)].astype(np.float32)
    kidx, nrm = io.sampling_surface_normal(plane, rng)
    assert abs(len(kidx) / 4000 - 0.5) < 0.05 and np.all(np.diff(kidx) > 0)
    n0 = np.array([-0.2, 0, 1.0]) / np.linalg.norm([-0.2, 0, 1.0])
    assert np.median(np.abs(nrm @ n0)) > 0.995


def test_oracle_icp_recovers_the_true_transform():
    from oracle import icp_oracle as io
    sc = make_scene(seed=0)
    e0 = pose_error(sc["T_init"], sc["T_true"])
    r = io.icp(sc["reading"], sc["map"], sc["T_init"], 50.0, seed=1)
    e1 = pose_error(r["T_LtoM"], sc["T_true"])
    assert e0[0] > 1.0 and e0[1] > 0.06
    assert r["success"] and 3 < r["iterations"] < 40
    assert e1[0] < 0.25 and e1[1] < 0.02, (e0, e1)


def test_oracle_icp_reports_failure_when_capped():
    from oracle import icp_oracle as io
    sc = make_scene(seed=1)
    r = io.icp(sc["reading"], sc["map"], sc["T_init"], 50.0, seed=1, max_iter=2)
    assert r["max_iter_reached"] and not r["success"] and np.array_equal(r["T_LtoM"], sc["T_init"])
