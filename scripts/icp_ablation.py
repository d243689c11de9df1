#!/usr/bin/env python
"""ICP pre-alignment on the MINS simulation's own geometry (UD_Warehouse), CPU only: does the device's restatement of the
libpointmatcher chain differ from the reference's where it deliberately deviates?

Variants of oracle/icp_oracle.py::icp over the same scenes and seeds:
  bins / unbounded   the reference: median-split bins of <= 7 points for the normals, unbounded 1-NN matcher
  bins / 2 m         the matcher bound alone
  knn7 / unbounded   the 7-NN normals alone
  knn7 / 2 m         what mins_b200/csrc/icp.cuh does
A scene = a pose on the UD_Warehouse B-spline trajectory, the local map a LiDAR would hold there (the dense 180 x 361 ground-truth
scan of the start pose and the 17 x 181 scans of the last second, voxel-thinned at 0.3 m like map_downsample_size), a new 17 x 181
scan with raw_noise, and an initial guess off by what the filter's prediction is typically off (0.5 deg, 3 cm).

  python scripts/icp_ablation.py [--seeds 64] > profiles/r2_icp_ablation.json
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mins_b200.sim.se3 import exp_so3, quat_jpl_to_rot  # noqa: E402
from mins_b200.sim.traj import load_bspline_dataset  # noqa: E402
from mins_b200.sim.world import PlaneWorld, lidar_ray_dirs, load_floorplan  # noqa: E402
from oracle import icp_oracle as io  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")
V_ANGLES = np.arange(-80, 81, 10.0)  # mins/config/simulation/config_lidar.yaml


def pose_error(T, T_true):
    dR = T[:3, :3].astype(np.float64) @ T_true[:3, :3].T
    return (float(np.degrees(np.arccos(np.clip((np.trace(dR) - 1) / 2, -1, 1)))),
            float(np.linalg.norm(T[:3, 3].astype(np.float64) - T_true[:3, 3])))


def make_scene(world, knots, seed, noise=0.01):
    rng = np.random.default_rng(1000 + seed)
    sparse = lidar_ray_dirs(V_ANGLES, -180, 180, 2.0)
    dense = lidar_ray_dirs(np.arange(-89, 90, 1.0), -180, 180, 1.0)

    def pose(k):
        R_ItoG = quat_jpl_to_rot(knots[k, 4:8]).T
        return R_ItoG, knots[k, 1:4]

    def scan(R, p, dirs):
        d = world.raycast(p, dirs @ R.T)
        keep = (d > 1.0) & (d < 100.0)
        pts = dirs[keep] * d[keep, None]
        return pts + noise * rng.standard_normal(pts.shape)

    # anchor = a knot in the moving part of the trajectory; the new scan is taken two knots (0.4 s) later
    k0 = int(rng.integers(30, len(knots) - 8))
    R_M, p_M = pose(k0)
    R_L, p_L = pose(k0 + 2)
    to_map = lambda pts, R, p: (pts @ R.T + p - p_M) @ R_M
    clouds = [scan(R_M, p_M, dense)]
    for dk in (-2, -1, 0, 1):
        R, p = pose(k0 + dk)
        clouds.append(to_map(scan(R, p, sparse), R, p))
    m = np.concatenate(clouds)
    key = np.floor(m / 0.3).astype(np.int64)
    _, first = np.unique(key, axis=0, return_index=True)
    m = m[np.sort(first)].astype(np.float32)
    T_true = np.eye(4)
    T_true[:3, :3] = R_M.T @ R_L
    T_true[:3, 3] = R_M.T @ (p_L - p_M)
    e = rng.standard_normal(3)
    T_init = np.eye(4)
    T_init[:3, :3] = exp_so3(np.deg2rad(0.5) * e / np.linalg.norm(e)) @ T_true[:3, :3]
    e = rng.standard_normal(3)
    T_init[:3, 3] = T_true[:3, 3] + 0.03 * e / np.linalg.norm(e)
    return dict(map=m, reading=scan(R_L, p_L, sparse).astype(np.float32), T_true=T_true, T_init=T_init.astype(np.float32))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=64)
    args = ap.parse_args()
    world = PlaneWorld(load_floorplan(os.path.join(G, "UD_Warehouse_floorplan.txt")))
    knots = load_bspline_dataset(os.path.join(G, "UD_Warehouse_bspline.txt"))
    variants = {"bins_unbounded": ("bins", None), "bins_2m": ("bins", 2.0), "knn7_unbounded": ("knn", None), "knn7_2m": ("knn", 2.0)}
    res = {k: [] for k in variants}
    init, beyond, maxm, trim = [], [], [], []
    for s in range(args.seeds):
        sc = make_scene(world, knots, s)
        init.append(pose_error(sc["T_init"], sc["T_true"]))
        for name, (nrm, bound) in variants.items():
            st = {} if name == "bins_unbounded" else None
            r = io.icp(sc["reading"], sc["map"], sc["T_init"], 50.0, seed=s, normals=nrm, max_match_dist=bound, stats=st)
            res[name].append(pose_error(r["T_icp"], sc["T_true"]) + (int(r["success"]), r["iterations"]))
            if st is not None:
                beyond.append(max(st["beyond_2m"]))
                maxm.append(max(st["max_match"]))
                trim.append(max(st["trim_limit"]))
        print(f"seed {s}: " + "  ".join(f"{k} {res[k][-1][0]:.3f} deg {res[k][-1][1] * 100:.1f} cm" for k in variants), file=sys.stderr)
    out = {"scenes": args.seeds, "initial_guess_error_deg_m": [float(np.mean([a for a, _ in init])), float(np.mean([b for _, b in init]))],
           "reference_chain_matches": {"max_match_dist_m": [float(np.mean(maxm)), float(np.max(maxm))],
                                       "reading_points_beyond_2m_max_over_iterations": [float(np.mean(beyond)), int(np.max(beyond))],
                                       "trimmed_filter_limit_m": [float(np.mean(trim)), float(np.max(trim))]}}
    for k, v in res.items():
        a = np.array(v, float)
        out[k] = {"rot_deg_mean_std_max": [float(a[:, 0].mean()), float(a[:, 0].std()), float(a[:, 0].max())],
                  "trans_m_mean_std_max": [float(a[:, 1].mean()), float(a[:, 1].std()), float(a[:, 1].max())],
                  "success": int(a[:, 2].sum()), "iterations_mean": float(a[:, 3].mean())}
    ref = np.array(res["bins_unbounded"], float)
    for k in ("bins_2m", "knn7_unbounded", "knn7_2m"):
        a = np.array(res[k], float)
        out[k]["paired_difference_to_reference_rot_deg_mean_std"] = [float((a[:, 0] - ref[:, 0]).mean()), float((a[:, 0] - ref[:, 0]).std())]
        out[k]["paired_difference_to_reference_trans_m_mean_std"] = [float((a[:, 1] - ref[:, 1]).mean()), float((a[:, 1] - ref[:, 1]).std())]
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
