#!/usr/bin/env python
"""Wall clock of LidarHelper::register_scan (ml_map_add_scan) and of the re-anchor on the G-64 workload: the 131 072-ray scan
into the 1 M-point map, repeated from the same starting map (the map is rebuilt, untimed, between repetitions).
MI_IT repetitions (default 20). Under `ncu --metrics gpu__time_duration.sum` the same script gives the per-kernel list."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from mins_b200 import synth
from mins_b200.lidar import LidarMap, OptionsLidar

op = OptionsLidar(raw_do_downsample=False)
w = synth.make_workload("G-64", op=op)
n = int(w["n_rays"])
m = LidarMap(max_scan_points=n, max_map_points=len(w["map"]) + 2 * n)
it = int(os.environ.get("MI_IT", "20"))
t_ins, t_re, t_build = [], [], []
T = np.eye(4)
T[:3, 3] = [0.05, -0.02, 0.01]
for r in range(it):
    t0 = time.perf_counter()
    m.Build(w["map"])
    t_build.append(time.perf_counter() - t0)
    m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    torch.cuda.synchronize()
    l0 = m.launch_count()
    t0 = time.perf_counter()
    n1 = m.register_scan(op)
    t_ins.append(time.perf_counter() - t0)
    l1 = m.launch_count()
    t0 = time.perf_counter()
    n2 = m.propagate_map_to_newest_clone(T, 1.0, op)
    t_re.append(time.perf_counter() - t0)
    l2 = m.launch_count()
print(f"map {len(w['map'])} -> {n1} after insert -> {n2} after re-anchor")
print(f"insert   us: min {min(t_ins) * 1e6:.1f} median {np.median(t_ins) * 1e6:.1f}  launches {l1 - l0}")
print(f"reanchor us: min {min(t_re) * 1e6:.1f} median {np.median(t_re) * 1e6:.1f}  launches {l2 - l1}")
print(f"build(1M, incl. 16 MB H2D) us: min {min(t_build) * 1e6:.1f} median {np.median(t_build) * 1e6:.1f}")
