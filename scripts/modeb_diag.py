import sys, os, time, ctypes as C
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from mins_b200 import synth, capi
from mins_b200.lidar import LidarMap, OptionsLidar
op = OptionsLidar(raw_do_downsample=True)
w = synth.make_workload("G-64", op=OptionsLidar(raw_do_downsample=False))
n = int(w["n_rays"])
m = LidarMap(max_scan_points=n, max_map_points=len(w["map"]) + 2 * n)
m.Build(w["map"])
for leaf in (0.3, 0.1):
    op.raw_downsample_size = leaf
    for _ in range(int(os.environ.get("MB_WARM", "200"))):
        rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    acc = {}
    for _ in range(int(os.environ.get("MB_IT", "50"))):
        rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
        for k, v in m.timing().items():
            acc[k] = acc.get(k, 0) + v / int(os.environ.get("MB_IT", "50"))
    print(leaf, cnt.n_points, {k: round(v * 1e3, 1) for k, v in acc.items()}, "launches/scan", (m.launch_count()) )
