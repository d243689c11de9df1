#!/usr/bin/env python
"""Per-CTA timeline of the kNN and rows kernels of one G-64 scan (ml_debug_timeline: %globaltimer stamps by thread 0 of every
CTA). Prints, per mark, min / median / max over the CTAs in microseconds after the first CTA of the kNN kernel started.

  kNN  marks (fused kernel): 0 entry, 1 after griddepcontrol.wait, 2 producer phase over, 3 exit; producers, last tile:
              4 loads landed + range test, 5 look-back done, 6 clouds written + hash filled, 7 scan + scatter + unit records
  rows marks: 0 entry, 1 after griddepcontrol.wait, 2 constants staged, 3 warp 0 left the tile loop, 4 CTA barrier passed,
              5 ticket taken, 6 (last CTA) partials summed + T staged, 7 (last CTA) result block written
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from mins_b200 import capi, synth  # noqa: E402
from mins_b200.lidar import LidarMap, OptionsLidar  # noqa: E402

shape = os.environ.get("TL_SHAPE", "G-64")
op = OptionsLidar(raw_do_downsample=False)
w = synth.make_workload(shape, op=op)
n = int(w["n_rays"])
m = LidarMap(max_scan_points=n, max_map_points=len(w["map"]) + 2 * n)
m.Build(w["map"])
dev = torch.from_numpy(w["raw"]).cuda()
pin = torch.from_numpy(w["raw"]).pin_memory()
src = pin if os.environ.get("TL_HOST") else dev
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
hx = w["total_hx"]
out = (np.zeros((hx, hx)), np.zeros(hx), capi.ml_counters())
T = np.ascontiguousarray(w["T_LtoM"], np.float32)
P = np.ascontiguousarray(w["P"], np.float64)


def step():
    rc = m.lib.ml_scan_update(m.h, src.data_ptr(), 4, n, w["t_rel"], op.min_range, op.max_range, 0.0, T.ctypes.data, C.byref(w["consts"]),
                              P.ctypes.data, out[0].ctypes.data, out[1].ctypes.data, C.byref(out[2]))
    assert rc == 0


m.set_timing(0)
for _ in range(20):
    step()
m.lib.ml_debug_timeline(m.h, 1, None, 0)
buf = np.zeros(2 * 2048 * 8, np.uint64)
for rep in range(3):
    if not os.environ.get("TL_NO_FLUSH"):
        flush.zero_()
    torch.cuda.synchronize()
    step()
    k = m.lib.ml_debug_timeline(m.h, 1, buf.ctypes.data, len(buf))
    a = buf.reshape(2, 2048, 8).astype(np.int64)
    t0 = a[0][a[0] > 0].min()
    print(f"--- scan {rep} ({shape}), microseconds after the first kNN CTA started")
    for name, arr in (("kNN ", a[0]), ("rows", a[1])):
        for s in range(8):
            v = arr[:, s]
            v = v[v > 0]
            if len(v) and not (name == "rows" and s == 7 and len(v) > 1):
                v = (v - t0) * 1e-3
                print(f"  {name} mark {s}: ctas {len(v):5d}  min {v.min():8.2f}  median {np.median(v):8.2f}  max {v.max():8.2f}")
    if rep == 2:  # tile-loop duration of every rows CTA against the SM it ran on
        r = a[1]
        live = (r[:, 2] > 0) & (r[:, 4] > 0)
        dur = (r[live, 4] - r[live, 2]) * 1e-3
        smid = np.where(r[live, 7] < 4096, r[live, 7], -1)
        order = np.argsort(smid)
        print("  rows tile loop (mark 2 -> 4) by SM id: " + " ".join(f"{int(sm)}:{d:.1f}" for sm, d in zip(smid[order], dur[order])))
m.close()
