#!/usr/bin/env python
"""Per-warp profile of the consumer phase of the fused front + kNN kernel (a -DML_KNN_PROF=1 build of the library, see fused.cuh):
how long every warp waited for tickets and tags, walked, emitted; how many units it got; when it started and ended.

  MINS_LIDAR_LIB=mins_b200/libmins_lidar_prof.so python scripts/knn_warp_profile.py
(build: nvcc <flags of __graft_entry__> -DML_KNN_PROF=1 mins_b200/csrc/api.cu -o mins_b200/libmins_lidar_prof.so)
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
PER = 8192 * 8
buf = np.zeros(2 * PER, np.uint64)
for rep in range(2):
    flush.zero_()
    torch.cuda.synchronize()
    step()
    k = m.lib.ml_debug_timeline(m.h, 1, buf.ctypes.data, len(buf))
    assert k == len(buf), k
    a = buf[:PER].reshape(-1, 8).astype(np.int64)
    a = a[(a[:, 5] > 0) | (a[:, 0] > 0)]
    rec = a[:, 3] >> 16
    a[:, 3] &= 0xffff
    inner = np.stack([a[:, 4] & 0xfffff, (a[:, 4] >> 20) & 0xfffff, (a[:, 4] >> 40) & 0xfffff], 1)
    a = np.concatenate([a, rec[:, None], inner], 1)
    rows = buf[PER:].reshape(-1, 8).astype(np.int64)
    got = a[a[:, 3] > 0]
    t0 = got[:, 5].min()
    us = lambda v: v * 1e-3
    q = lambda v: " ".join(f"{x:7.2f}" for x in np.percentile(v, [0, 10, 50, 90, 100]))
    print(f"--- scan {rep} ({shape}): {len(a)} warps reported, {len(got)} got units, {int(a[:, 3].sum())} units")
    print(f"  units per warp        : {q(got[:, 3])}   (min p10 p50 p90 max)")
    print(f"  first unit start [us] : {q(us(got[:, 5] - t0))}")
    print(f"  last unit end    [us] : {q(us(got[:, 6] - t0))}")
    print(f"  wait per warp    [us] : {q(us(got[:, 0]))}")
    print(f"  walk per warp    [us] : {q(us(got[:, 1]))}")
    print(f"  emit per warp    [us] : {q(us(got[:, 2]))}")
    print(f"  walk per unit    [us] : {q(us(got[:, 1]) / got[:, 3])}")
    print(f"  emit per unit    [us] : {q(us(got[:, 2]) / got[:, 3])}")
    print(f"  record+query per unit : {q(us(got[:, 8]) / got[:, 3])}")
    print(f"  rows set-up per unit  : {q(us(got[:, 9]) / got[:, 3])}")
    print(f"  stage wait per unit   : {q(us(got[:, 10]) / got[:, 3])}")
    print(f"  test per unit         : {q(us(got[:, 11]) / got[:, 3])}")
    for role in (0, 1):
        g = got[got[:, 7] == role]
        if len(g):
            print(f"  {'producer' if role else 'consumer'} CTAs: warps {len(g)}, units {int(g[:, 3].sum())}, walk/unit median {np.median(us(g[:, 1]) / g[:, 3]):.2f} us, "
                  f"last end median {np.median(us(g[:, 6] - t0)):.2f} us")
    rs = rows[rows[:, 0] > 0]
    if len(rs):
        print(f"  rows kernel first CTA entry at {us(rs[:, 0].min() - t0):.2f} us, exit max {us(rs[:, 5].max() - t0):.2f} us")
m.close()
