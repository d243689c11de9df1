"""GPU-box diagnostic: per-iteration wall time of the pinned-host ml_scan_update over the life of a process, with the PCIe
link generation/width (NVML), the CPU the thread runs on and the NUMA layout — to explain a bimodal H2D cost."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import pynvml as nv  # noqa: E402
from mins_b200 import capi, synth  # noqa: E402
from mins_b200.lidar import LidarMap, OptionsLidar  # noqa: E402

nv.nvmlInit()
hd = nv.nvmlDeviceGetHandleByIndex(0)
libc = C.CDLL("libc.so.6")
os.system("lscpu | grep -i -E 'numa|socket|model name' ; nvidia-smi topo -m 2>/dev/null | head -8; cat /sys/bus/pci/devices/*/numa_node 2>/dev/null | sort | uniq -c")
print("affinity:", len(os.sched_getaffinity(0)), "cpus")
op = OptionsLidar(raw_do_downsample=False)
w = synth.make_workload("G-64", op=op)
n, hx = int(w["n_rays"]), w["total_hx"]
m = LidarMap(max_scan_points=n, max_map_points=len(w["map"]) + 2 * n)
m.Build(w["map"])
pinned = torch.empty((n, 4), dtype=torch.float32).pin_memory()
pinned.copy_(torch.from_numpy(w["raw"]))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
out = (np.zeros((hx, hx)), np.zeros(hx), capi.ml_counters())
T = np.ascontiguousarray(w["T_LtoM"], np.float32)
P = np.ascontiguousarray(w["P"], np.float64)
lib, h, cns = m.lib, m.h, w["consts"]
ptr = pinned.numpy().ctypes.data
m.set_timing(0)


def step():
    rc = lib.ml_scan_update(h, ptr, 4, n, w["t_rel"], op.min_range, op.max_range, 0.0, T.ctypes.data, C.byref(cns), P.ctypes.data,
                            out[0].ctypes.data, out[1].ctypes.data, C.byref(out[2]))
    assert rc == 0, rc


def link():
    return (nv.nvmlDeviceGetCurrPcieLinkGeneration(hd), nv.nvmlDeviceGetCurrPcieLinkWidth(hd), nv.nvmlDeviceGetMaxPcieLinkGeneration(hd),
            nv.nvmlDeviceGetPerformanceState(hd), nv.nvmlDeviceGetClockInfo(hd, nv.NVML_CLOCK_MEM), nv.nvmlDeviceGetClockInfo(hd, nv.NVML_CLOCK_SM))


t_start = time.perf_counter()
for mode in ("flush", "noflush", "flush", "sleep1ms+flush", "flush"):
    for blk in range(6):
        ts = []
        for _ in range(100):
            if "flush" in mode and mode != "noflush":
                flush.zero_()
            torch.cuda.synchronize()
            if mode.startswith("sleep"):
                time.sleep(0.001)
            t0 = time.perf_counter()
            step()
            ts.append(time.perf_counter() - t0)
        ts = np.array(ts) * 1e6
        print(f"t={time.perf_counter() - t_start:6.2f}s {mode:15s} median {np.median(ts):6.1f} p10 {np.percentile(ts, 10):6.1f} p90 {np.percentile(ts, 90):6.1f} "
              f"cpu {libc.sched_getcpu()} link(gen,width,maxgen,pstate,memclk,smclk) {link()}", flush=True)
# raw H2D timing with events
d = torch.empty((n, 4), dtype=torch.float32, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for rep in range(5):
    tt = []
    for _ in range(50):
        flush.zero_()
        torch.cuda.synchronize()
        e0.record()
        d.copy_(pinned, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        tt.append(e0.elapsed_time(e1) * 1e3)
    print(f"H2D 2 MB device-timed after flush: median {np.median(tt):.1f} us p10 {np.percentile(tt, 10):.1f} p90 {np.percentile(tt, 90):.1f}", link())
m.close()
