#!/usr/bin/env python
"""Closed loop on UD_Warehouse with the CPU ORACLE doing every LiDAR update (no GPU): map_use_icp off, on with the reference's chain
(median-split bins, unbounded matcher) and on with the device's variant (7-NN normals, matcher bounded at 2 m). Answers in the loop
what scripts/icp_ablation.py answers per alignment: is the loss with ICP on a property of the method or of the device's deviations?

  python scripts/icp_closed_loop_ablation.py [--seeds 2] [--duration 10] > profiles/r2_icp_closed_loop_ablation.json
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mins_b200.lidar import OptionsLidar  # noqa: E402
from mins_b200.sim.runner import SimConfig, run  # noqa: E402
from oracle.backend import OracleBackend  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=2)
    ap.add_argument("--duration", type=float, default=10.0)
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--variant", default="", help="run one variant only (the three are independent: one process each)")
    args = ap.parse_args()
    variants = {"icp_off": None, "icp_reference_chain": ("bins", None), "icp_device_variant": ("knn", 2.0)}
    if args.variant:
        variants = {args.variant: variants[args.variant]}
    out = {"duration_s": args.duration, "seeds": args.seeds}
    for name, v in variants.items():
        rows = []
        for seed in range(args.seeds):
            op = OptionsLidar(raw_do_downsample=True, map_use_icp=v is not None)
            cfg = SimConfig(seed=seed, duration=args.duration, bspline_path=os.path.join(G, "UD_Warehouse_bspline.txt"),
                            floorplan_path=os.path.join(G, "UD_Warehouse_floorplan.txt"), use_gt_lidar=True, op=op)
            be = OracleBackend(threads=args.threads) if v is None else OracleBackend(threads=args.threads, icp_normals=v[0], icp_max_match_dist=v[1])
            t0 = time.time()
            r = run(cfg, be)
            rows.append([r["rmse_ori_deg"], r["rmse_pos_m"], r["nees_pos"], be.icp_runs, be.icp_ok])
            print(f"{name} seed {seed}: {r['rmse_ori_deg']:.3f} deg {r['rmse_pos_m'] * 100:.1f} cm NEES {r['nees_pos']:.2f} "
                  f"icp {be.icp_ok}/{be.icp_runs} ({time.time() - t0:.0f} s)", file=sys.stderr)
        a = np.array(rows, float)
        out[name] = {"rmse_ori_deg_mean_std": [float(a[:, 0].mean()), float(a[:, 0].std())],
                     "rmse_pos_m_mean_std": [float(a[:, 1].mean()), float(a[:, 1].std())],
                     "nees_pos_mean": float(a[:, 2].mean()), "icp_success": [int(a[:, 4].sum()), int(a[:, 3].sum())], "per_seed": rows}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
