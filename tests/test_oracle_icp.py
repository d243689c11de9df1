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


def test_device_variants_of_the_chain_do_not_degrade_the_alignment():
    """The device chain (mins_b200/csrc/icp.cuh) deviates from libpointmatcher in two places: normals from the 7 nearest map points
    instead of median-split bins, and a matcher bounded at 2 m. On UD_Warehouse scenes (scripts/icp_ablation.py, 64 of them in
    profiles/r2_icp_ablation.md) the bound changes nothing - the trimmed-distance filter cuts far below it - and the 7-NN normals
    align at least as well as the bins."""
    import importlib.util
    import os
    from conftest import ROOT
    from oracle import icp_oracle as io
    spec = importlib.util.spec_from_file_location("icp_ablation", os.path.join(ROOT, "scripts", "icp_ablation.py"))
    ab = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ab)
    world = ab.PlaneWorld(ab.load_floorplan(os.path.join(ab.G, "UD_Warehouse_floorplan.txt")))
    knots = ab.load_bspline_dataset(os.path.join(ab.G, "UD_Warehouse_bspline.txt"))
    for seed in (0, 4):
        sc = ab.make_scene(world, knots, seed)
        e0 = ab.pose_error(sc["T_init"], sc["T_true"])
        st = {}
        ref = io.icp(sc["reading"], sc["map"], sc["T_init"], 50.0, seed=seed, stats=st)
        bound = io.icp(sc["reading"], sc["map"], sc["T_init"], 50.0, seed=seed, max_match_dist=2.0)
        dev = io.icp(sc["reading"], sc["map"], sc["T_init"], 50.0, seed=seed, normals="knn", max_match_dist=2.0)
        e_ref, e_bound, e_dev = (ab.pose_error(r["T_icp"], sc["T_true"]) for r in (ref, bound, dev))
        assert ref["success"] and bound["success"] and dev["success"]
        assert e_ref[0] < e0[0] and e_dev[0] < e0[0]                      # both improve on the 0.5 deg / 3 cm guess in rotation
        assert max(st["trim_limit"]) < 1.0 < 2.0                          # the 85 % quantile of the match distances is far below 2 m
        assert abs(e_bound[0] - e_ref[0]) < 0.08 and abs(e_bound[1] - e_ref[1]) < 0.01
        assert e_dev[0] < e_ref[0] + 0.1 and e_dev[1] < e_ref[1] + 0.02
