"""GPU-box diagnostic (not a test): where the end-to-end time of ml_scan_update goes on G-64.
Wall clock of the blocking C-ABI call for device-resident vs pinned-host vs pageable-host input, with and without the
L2 flush and the stage timers; plus the raw cost of the 2 MB H2D copy and of an empty launch + synchronise."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from mins_b200 import capi, synth  # noqa: E402
from mins_b200.lidar import LidarMap, OptionsLidar  # noqa: E402

op = OptionsLidar(raw_do_downsample=False)
w = synth.make_workload(sys.argv[1] if len(sys.argv) > 1 else "G-64", op=op)
n, hx = int(w["n_rays"]), w["total_hx"]
m = LidarMap(max_scan_points=n, max_map_points=len(w["map"]) + 2 * n)
m.Build(w["map"])
pinned = torch.empty((n, 4), dtype=torch.float32).pin_memory()
pinned.copy_(torch.from_numpy(w["raw"]))
raw_dev = torch.from_numpy(w["raw"]).cuda()
pageable = np.ascontiguousarray(w["raw"])
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
out = (np.zeros((hx, hx)), np.zeros(hx), capi.ml_counters())
T = np.ascontiguousarray(w["T_LtoM"], np.float32)
P = np.ascontiguousarray(w["P"], np.float64)
lib, h, cns = m.lib, m.h, w["consts"]


def step(ptr):
    rc = lib.ml_scan_update(h, ptr, 4, n, w["t_rel"], op.min_range, op.max_range, 0.0, T.ctypes.data, C.byref(cns), P.ctypes.data,
                            out[0].ctypes.data, out[1].ctypes.data, C.byref(out[2]))
    assert rc == 0, rc


def bench(name, ptr, do_flush, reps=200):
    for _ in range(10):
        step(ptr)
    ts = []
    for _ in range(reps):
        if do_flush:
            flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        step(ptr)
        ts.append(time.perf_counter() - t0)
    ts = np.array(ts) * 1e6
    print(f"{name:55s} median {np.median(ts):7.1f} us  p10 {np.percentile(ts, 10):7.1f}  p90 {np.percentile(ts, 90):7.1f}", flush=True)


for timing in (1, 0):
    m.set_timing(timing)
    for fl in (True, False):
        tag = f"timers={'on' if timing else 'off'} flush={'yes' if fl else 'no'}"
        bench(f"device-resident input   {tag}", raw_dev.data_ptr(), fl)
        bench(f"pinned host input       {tag}", pinned.numpy().ctypes.data, fl)
        bench(f"pageable host input     {tag}", pageable.ctypes.data, fl)
m.set_timing(1)
step(raw_dev.data_ptr())
print("stage ms (device events):", {k: round(v, 4) for k, v in m.timing().items()})

# raw copies / launch latency with torch on its own stream
d = torch.empty((n, 4), dtype=torch.float32, device="cuda")
for sz in (n, n // 2, n // 4, n // 8):
    ts = []
    for _ in range(200):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        d[:sz].copy_(pinned[:sz], non_blocking=True)
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    print(f"H2D {sz * 16 / 1e6:.2f} MB pinned + sync: median {np.median(ts) * 1e6:.1f} us -> {sz * 16 / np.median(ts) / 1e9:.1f} GB/s")
ts = []
z = torch.zeros(1, device="cuda")
for _ in range(200):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    z.add_(1)
    torch.cuda.synchronize()
    ts.append(time.perf_counter() - t0)
print(f"one tiny kernel + sync: median {np.median(ts) * 1e6:.1f} us")
hres = torch.empty(33368, dtype=torch.uint8).pin_memory()
dres = torch.empty(33368, dtype=torch.uint8, device="cuda")
ts = []
for _ in range(200):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    hres.copy_(dres, non_blocking=True)
    torch.cuda.synchronize()
    ts.append(time.perf_counter() - t0)
print(f"D2H 33 KB + sync: median {np.median(ts) * 1e6:.1f} us")
m.close()
