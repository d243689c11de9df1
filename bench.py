#!/usr/bin/env python
"""bench.py — LiDAR scan points registered per second on B200 (BASELINE.json metric), one JSON line.

A "step" = one LiDAR scan through the hot path (range filter -> transform -> exact kNN vs the local map -> plane fit ->
residual/Jacobian -> chi2 gate -> reduction to (HtH, Htr) -> D2H of the small system), against the resident map.
Workload (N=1): BASELINE.json configs[2] "synthetic 64-beam scans (~130k pts/scan) vs 1M-point voxel map, k=5" — the
configuration the metric's target is quoted on; it fits one GPU.

  value : whole-job points/s with the raw scan already resident in HBM (device-pointer input), device time from CUDA
          events on the handle's own stream (ml_last_timing), L2 flushed between steps.
  e2e   : the same through the C-ABI call with PINNED HOST input and HOST output inside the timed region (wall clock of the
          blocking call). The bytes cross PCIe inside the call: every CTA of the scan kernel reads its tile of the pinned scan
          directly (TMA bulk copy, each point once) and writes its reduced system into pinned host memory, where the library's
          host tail adds the parts as they land.
  N > 1 : independent Monte-Carlo replicas, one process per GPU, no data-path collective: weak scaling.
  --impl reference : the reference CPU path on the host cores (reference ikd-Tree compiled unmodified + restated rows).
"""
import argparse
import gc
import ctypes as C
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "LiDAR scan points registered/sec per B200"
UNIT = "points/s"
WORKLOAD = "G-64"
ALG_BYTES_PER_POINT = 96  # 16 B scan point + 5 x 16 B neighbours
CPU_BUDGET_S = 300.0  # the reference arm must end within a few minutes whatever K and W are asked for


L2_NOTE = ("GPU arm: flushed (256 MiB write) between steps; reference arm: host caches, not flushed (map 16 MB + tree 160 MB "
           "exceed them)")


def workload_desc(w, n_gpus, points_per_step, n_acc):
    """`config` of the JSON line: the same dict, key for key, on the GPU arm and on the reference arm (both run the same
    scan against the same map; both count the same accepted points)."""
    return {"workload": f"{WORKLOAD}: synthetic 64-beam x 2048-azimuth scan ({w['n_rays']} rays/scan) vs {len(w['map'])}-point "
                        f"local map, k=5, total_hx={w['total_hx']}, raw_do_downsample=false (every in-range ray is a query)",
            "points_per_step": int(points_per_step), "map_points": int(len(w["map"])), "seed": int(w["seed"]),
            "parallelism": f"{n_gpus} independent replica(s), no collective", "l2": L2_NOTE,
            "accepted_points_per_step": int(n_acc)}


class ClockSampler:
    """SM clock + throttle reasons read through NVML from the benchmark's own thread, once per step right after the step's
    blocking call returns (inside the timed region, outside the per-step timers). A background polling thread was measured to
    inflate the wall-clock e2e of a ~180 us call by ~140 us (GIL + NVML/driver lock contention), so there is none."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index):
        self.samples, self.err, self.sm_max, self.hdl, self.nv = [], None, None, None, None
        try:
            import pynvml as nv
            nv.nvmlInit()
            # LOCAL_RANK indexes CUDA_VISIBLE_DEVICES; map it back to the NVML index when the variable is set
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = index
            if vis:
                try:
                    idx = int(vis.split(",")[index])
                except Exception:
                    pass
            self.nv = nv
            self.hdl = nv.nvmlDeviceGetHandleByIndex(idx)
            self.sm_max = float(nv.nvmlDeviceGetMaxClockInfo(self.hdl, nv.NVML_CLOCK_SM))
        except Exception as e:  # pragma: no cover
            self.err = repr(e)

    def sample(self):
        if self.hdl is None:
            return
        try:
            nv = self.nv
            clk = nv.nvmlDeviceGetClockInfo(self.hdl, nv.NVML_CLOCK_SM)
            try:
                rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.hdl)
            except Exception:
                rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.hdl)
            self.samples.append((float(clk), int(rs)))
        except Exception as e:  # pragma: no cover
            self.err = repr(e)

    def finish(self):
        sm = [c for c, _ in self.samples]
        reasons = set()
        for _, r in self.samples:
            for bit, name in self.REASONS.items():
                if r & bit:
                    reasons.add(name)
        out = {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.sm_max, "reasons": sorted(reasons),
               "samples": len(sm), "how": "NVML, one sample per step right after the step's call returns, timed regions only"}
        if self.err:
            out["error"] = self.err
        return out


def bind_to_gpu_cpus(index):
    """one process per GPU, kept on the host cores next to that GPU (NVML's cpu affinity = the GPU's NUMA node): the pinned
    scan buffer and the pinned result block are then first-touched on that node, and the thread that waits for the result is
    not migrated across sockets. On this pool's boxes NVML lists every core for every GPU (one NUMA node): a no-op there,
    see profiles/r1_multi_gpu.md. Returns a description."""
    if os.environ.get("MINS_BENCH_NO_BIND"):
        return "unchanged (MINS_BENCH_NO_BIND)"
    try:
        import pynvml as nv
        nv.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = index
        if vis:
            try:
                idx = int(vis.split(",")[index])
            except Exception:
                pass
        hdl = nv.nvmlDeviceGetHandleByIndex(idx)
        n_cpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(hdl, (n_cpu + 63) // 64)
        cpus = {64 * wi + b for wi, wd in enumerate(words) for b in range(64) if (int(wd) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        want = cpus & allowed
        if not want or want == allowed:
            return f"unchanged ({len(allowed)} cpus allowed; NVML lists {len(cpus)} for GPU {idx})"
        os.sched_setaffinity(0, want)
        return f"{len(want)} of {len(allowed)} cpus (NVML cpu affinity of GPU {idx})"
    except Exception as e:  # pragma: no cover
        return f"unchanged ({e!r})"


def cpu_reference_run(w, steps, warmup, threads):
    """The reference CPU path for this workload: one step = the WHOLE scan through LiDARData ctor -> transform -> the
    reference's own ikd-Tree kNN in its static-chunk thread loop -> rows -> stacking -> Givens compression, i.e. exactly what
    UpdaterLidar::update does between the scan and EKFUpdate (UpdaterLidar.cpp:266-397). HtH / Htr are NOT formed (the
    reference never forms them). Steps and warm-up are run as asked; only if (K + W) whole scans would not fit CPU_BUDGET_S
    is the scan thinned to every s-th ray (a strided sample of all azimuths and rings, said in `sample`).
    Returns (points/s, ms/step, rays per step, description)."""
    from oracle import pyoracle as po
    if not po.ref_available():
        raise RuntimeError("oracle/_ref/libikd_ref.so missing (run __graft_entry__.build() in the build container)")
    tree = po.RefIkd(w["op"].map_downsample_size)
    t0 = time.perf_counter()
    tree.build(w["map"])
    t_build = time.perf_counter() - t0
    raw = np.ascontiguousarray(w["raw"])

    n_acc = [0]

    def one(r):
        t0 = time.perf_counter()
        res = tree.lidar_update(r, w["t_rel"], w["op"], w["T_LtoM"], w["consts"], w["P"], threads=threads, want_info=False)
        dt_ = time.perf_counter() - t0
        n_acc[0] = int(res[2].n_acc)
        return dt_

    t_probe = one(raw)  # untimed probe (also first-touch warm-up of the tree)
    stride = 1
    while stride < 64 and (steps + warmup) * t_probe / stride > CPU_BUDGET_S:
        stride *= 2
    if stride > 1:
        raw = np.ascontiguousarray(raw[::stride])
    ts = [one(raw) for _ in range(warmup + steps)][warmup:]
    per = float(np.mean(ts))
    what = "the whole scan" if stride == 1 else f"every {stride}-th ray of the scan (all azimuths and rings; {steps}+{warmup} whole scans exceed {CPU_BUDGET_S:.0f} s)"
    sample = (f"{what}: {len(raw)} of {w['n_rays']} rays per step vs the full {len(w['map'])}-point map; "
              f"kNN = reference mins/src/update/lidar/ikd_Tree.cpp compiled unmodified (Build {t_build:.2f} s, not timed), "
              f"rows/stacking/Givens compression = restated C++ (oracle/lidar_oracle.cpp), no HtH/Htr; reference threading "
              f"(static chunks, {threads} threads), -O3 no -march")
    return len(raw) / per, per * 1e3, len(raw), sample, n_acc[0]


def aux_measurements(m, w, op, step, host_ptr, out, flush_l2, hx, reps=10):
    """SURVEY.md §8(d) items (2) and (3) plus the ICP pre-alignment, wall clock of the blocking calls, outside the timed
    regions: ms per LiDAR EKF update (e2e step + host re-factorisation + EKF update of a 100-dim state through the 48
    columns), map-insert points/s (register_scan of the 131 072-point scan into the 1 M-point map), re-anchor ms, ICP ms."""
    import torch
    from mins_b200.lidar import info_to_sqrt
    rng = np.random.default_rng(0)
    L = rng.standard_normal((100, 100))
    Pfull = 1e-4 * (L @ L.T / 100 + np.eye(100))
    cols = np.arange(hx)
    t_upd = []
    for _ in range(reps):
        flush_l2()
        t0 = time.perf_counter()
        step(host_ptr)
        Hc, rc = info_to_sqrt(out[0], out[1])               # stand-in for measurement_compress_inplace (StateHelper.cpp:631-643)
        PHt = Pfull[:, cols] @ Hc.T                          # StateHelper::EKFUpdate (StateHelper.cpp:121-200), R = I
        S = Hc @ PHt[cols] + np.eye(len(rc))
        K = np.linalg.solve(S, PHt.T).T
        dx = K @ rc
        Pn = Pfull - K @ PHt.T
        t_upd.append(time.perf_counter() - t0)
        assert np.isfinite(dx).all() and np.isfinite(Pn).all()
    aux = {"ms_per_lidar_ekf_update": float(np.median(t_upd) * 1e3), "ekf_state_dim": 100, "total_hx": int(hx)}
    # mode B of SURVEY.md 8(d): raw_do_downsample with leaf 0.3 (PCL VoxelGrid restatement on the device before the kNN)
    import ctypes as C
    T = np.ascontiguousarray(w["T_LtoM"], np.float32)
    P = np.ascontiguousarray(w["P"], np.float64)
    t_b = []
    for _ in range(reps):
        flush_l2()
        t0 = time.perf_counter()
        rc = m.lib.ml_scan_update(m.h, host_ptr, 4, int(w["n_rays"]), w["t_rel"], op.min_range, op.max_range, 0.3, T.ctypes.data,
                                  C.byref(w["consts"]), P.ctypes.data, out[0].ctypes.data, out[1].ctypes.data, C.byref(out[2]))
        t_b.append(time.perf_counter() - t0)
        assert rc in (0, 1)
    aux["mode_B_downsample_leaf_0.3"] = {"e2e_ms_per_scan": float(np.median(t_b) * 1e3), "points_after_downsample": int(out[2].n_points),
                                         "raw_points_per_s": float(w["n_rays"] / np.median(t_b))}
    step(host_ptr)  # back to mode A: the resident scan is the full one again
    # ICP pre-alignment of the resident scan against the 1 M-point map (map_icp_dist 50 m)
    t_icp, iters = [], []
    for r in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _, res = m.icp(w["T_LtoM"], 50.0, seed=r)
        t_icp.append(time.perf_counter() - t0)
        iters.append(res.iterations)
    aux["icp_ms"] = float(np.median(t_icp) * 1e3)
    aux["icp_iterations"] = int(np.median(iters))
    aux["icp_points"] = {"reading": int(res.n_reading), "reference": int(res.n_reference)}
    # map insert (LidarHelper::register_scan) and re-anchor (propagate_map_to_newest_clone); both change the map: last.
    # The first insert adds the scan's voxels; the repetitions run the same contest again (same work, map unchanged).
    n0 = m.size()
    t_ins, t_defer, t_pair = [], [], []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n1 = m.register_scan(op)
        t_ins.append(time.perf_counter() - t0)
    for _ in range(reps):                                   # n_live == NULL: queued, the call returns at once
        m.size()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m.register_scan(op, wait=False)
        t_defer.append(time.perf_counter() - t0)
    m.size()
    for _ in range(reps):                                   # the estimator's per-scan pair, back to back (no host work between)
        t0 = time.perf_counter()
        step(host_ptr)
        m.register_scan(op, wait=False)
        t_pair.append(time.perf_counter() - t0)
    m.size()
    t_med = float(np.median(t_ins))
    aux["map_insert_points_per_s"] = float(w["n_rays"] / t_med)
    aux["map_insert_ms"] = t_med * 1e3
    aux["map_insert_deferred_call_ms"] = float(np.median(t_defer) * 1e3)
    aux["scan_update_plus_deferred_insert_ms"] = float(np.median(t_pair) * 1e3)
    aux["map_points_before_after_insert"] = [int(n0), int(n1)]
    T = np.eye(4)
    T[:3, 3] = [0.05, -0.02, 0.01]
    t0 = time.perf_counter()
    n2 = m.propagate_map_to_newest_clone(T, 1.0, op)
    aux["reanchor_ms"] = float((time.perf_counter() - t0) * 1e3)
    aux["map_points_after_reanchor"] = int(n2)
    return aux


def aux_config1(device, reps=20):
    """BASELINE.json configs[1]: 2 x VLP-16 (32 000 rays) against the KAIST-shaped map, for the simulation option set and the
    KAIST option set (mins/config/kaist/config_lidar.yaml:11-29: leaf 2.0, map voxel 0.5, map_ngbr_max_d 10 -> the cube walk
    reaches 7 cells), each with and without raw down-sampling. e2e = wall clock of ml_scan_update from pinned host memory,
    device = CUDA events around the whole step. knn_fail_share = queries get_neighbors rejects (for the KAIST radius these are
    the ones that walked to the last level)."""
    import torch
    from mins_b200 import capi, synth
    from mins_b200.lidar import LidarMap, OptionsLidar, kaist_options
    res = {}
    for name, mk, radius in (("sim_options", lambda ds: OptionsLidar(raw_do_downsample=ds), None),
                             ("kaist_options", lambda ds: kaist_options(raw_do_downsample=ds), 30.0)):
        for ds in (True, False):
            op = mk(ds)
            w = synth.make_workload("G-vlp", op=op, map_radius=radius)
            if name == "kaist_options":  # sparse map ahead of the sensor: part of the scan searches to the last level
                mp = w["map"]
                w["map"] = np.ascontiguousarray(mp[(mp[:, 0] < 3.0) | (np.arange(len(mp)) % 40 == 0)])
            n, hx = int(w["n_rays"]), w["total_hx"]
            m = LidarMap(max_scan_points=n, max_map_points=len(w["map"]) + 2 * n, device=device)
            m.Build(w["map"])
            pinned = torch.from_numpy(w["raw"]).pin_memory()
            out = (np.zeros((hx, hx)), np.zeros(hx), capi.ml_counters())
            T = np.ascontiguousarray(w["T_LtoM"], np.float32)
            P = np.ascontiguousarray(w["P"], np.float64)
            leaf = op.raw_downsample_size if ds else 0.0
            m.set_timing(2)
            tw, td = [], []
            for it in range(5 + reps):
                t0 = time.perf_counter()
                rc = m.lib.ml_scan_update(m.h, pinned.data_ptr(), 4, n, w["t_rel"], op.min_range, op.max_range, leaf, T.ctypes.data,
                                          C.byref(w["consts"]), P.ctypes.data, out[0].ctypes.data, out[1].ctypes.data, C.byref(out[2]))
                dt_ = time.perf_counter() - t0
                assert rc in (0, 1)
                if it >= 5:
                    tw.append(dt_)
                    td.append(m.timing()["total"])
            c = out[2]
            res[f"{name}_{'mode_B_leaf_%.1f' % leaf if ds else 'mode_A'}"] = {
                "e2e_ms_per_scan": float(np.median(tw) * 1e3), "device_ms_per_scan": float(np.median(td)), "rays": n,
                "map_points": int(len(w["map"])), "queries": int(c.n_points), "accepted": int(c.n_acc),
                "knn_fail_share": float(c.n_knn_fail / max(c.n_points, 1)), "raw_points_per_s_e2e": float(n / np.median(tw))}
            m.close()
    return res


def aux_monte_carlo(runs=8, per_gpu=8, duration=20.0):
    """BASELINE.json configs[4] in small: independent closed-loop runs of the host-C++ harness on this GPU (scripts/mc_sim_cpp.py)"""
    p = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "mc_sim_cpp.py"), "--gpus", "1", "--runs", str(runs), "--per-gpu",
                        str(per_gpu), "--duration", str(duration)], capture_output=True, text=True, timeout=600)
    if p.returncode != 0:
        return {"error": p.stderr[-300:]}
    return json.loads(p.stdout.strip().splitlines()[-1])


def emit(line):
    """the one JSON line, on the process's ORIGINAL stdout (see main: fd 1 is pointed at stderr while the run lasts)"""
    os.write(_STDOUT_FD, (json.dumps(line) + "\n").encode())


_STDOUT_FD = 1


def main():
    # stdout carries exactly one JSON line. Libraries write there too (NCCL prints "NCCL version ..." on stdout when the first
    # communicator comes up): keep a private copy of fd 1 for the line and send everything else to stderr.
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--warmup-seconds", type=float, default=0.5, help="minimum duration of the warm-up (0 for profiler runs)")
    ap.add_argument("--no-aux", action="store_true", help="skip the informational aux block (profiler runs)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    threads = os.cpu_count() or 1

    from mins_b200 import synth
    from mins_b200.lidar import OptionsLidar

    if args.impl == "reference":
        if rank != 0:
            return 0
        op = OptionsLidar(raw_do_downsample=False)
        w = synth.make_workload(WORKLOAD, op=op)
        steps, warmup = max(1, args.steps), max(0, args.warmup)
        val, ms, n_step, sample, n_acc = cpu_reference_run(w, steps, warmup, threads)
        line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
                "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32 kNN / f64 rows", "data": "synthetic",
                "config": workload_desc(w, args.gpus, n_step, n_acc),
                "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "reference", "sample": sample},
                "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return 0

    affinity = bind_to_gpu_cpus(local_rank)
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the LiDAR path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from mins_b200 import capi
    from mins_b200.lidar import LidarMap

    op = OptionsLidar(raw_do_downsample=False)
    # every rank = an independent replica with its own world/scan (Monte-Carlo unit)
    # every rank runs the SAME synthetic world (weak scaling = fixed work per GPU): with a world of its own per rank the max over
    # ranks measured which world was the heaviest (58.8 us at N = 1 against 69 us at N = 8 for identical code), not the hardware
    w = synth.make_workload(WORKLOAD, op=op, seed=64)
    n = int(w["n_rays"])
    hx = w["total_hx"]
    m = LidarMap(max_scan_points=n, max_map_points=len(w["map"]) + 2 * n, device=local_rank,
                 grid_cell=float(os.environ.get("MINS_GRID_CELL", "0")))  # 0 = library default (0.5 m); env = tuning experiments
    m.Build(w["map"])
    pinned = torch.empty((n, 4), dtype=torch.float32).pin_memory()
    pinned.copy_(torch.from_numpy(w["raw"]))
    raw_pinned = pinned.numpy()
    raw_dev = torch.from_numpy(w["raw"]).cuda()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    out = (np.zeros((hx, hx)), np.zeros(hx), capi.ml_counters())
    lib, h = m.lib, m.h
    T = np.ascontiguousarray(w["T_LtoM"], np.float32)
    P = np.ascontiguousarray(w["P"], np.float64)
    cns = w["consts"]

    def step(ptr):
        rc = lib.ml_scan_update(h, ptr, 4, n, w["t_rel"], op.min_range, op.max_range, 0.0, T.ctypes.data, C.byref(cns), P.ctypes.data,
                                out[0].ctypes.data, out[1].ctypes.data, C.byref(out[2]))
        if rc != 0:
            raise RuntimeError(f"ml_scan_update -> {rc}: {lib.ml_last_error(h).decode()}")

    def flush_l2():
        flush.zero_()
        torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    dev_ptr, host_ptr = raw_dev.data_ptr(), raw_pinned.ctypes.data
    # warm-up: at least W steps of both input kinds AND at least 0.5 s of sustained work. Measured on this pool
    # (scripts/h2d_hugepage_diag.py): for the first ~50-100 ms of PCIe activity of a process a 2 MB pinned H2D copy takes
    # ~180 us instead of 46 us (link / DMA path still ramping up); the metric is defined at steady state.
    t_w = time.perf_counter()
    n_w = 0
    while n_w < max(3, args.warmup) or time.perf_counter() - t_w < args.warmup_seconds:
        flush_l2()
        step(dev_ptr)
        step(host_ptr)
        n_w += 1
    sampler = ClockSampler(local_rank)
    # ---- timed region 1: device-resident input, device time (two CUDA events on the handle's stream around the whole step) -
    m.set_timing(2)
    barrier()
    l0 = m.launch_count()
    t_wall0 = time.perf_counter()
    dev_ms = 0.0
    for _ in range(args.steps):
        flush_l2()
        step(dev_ptr)
        sampler.sample()
        dev_ms += m.timing()["total"]
    barrier()
    wall_total = time.perf_counter() - t_wall0
    launches = m.launch_count() - l0
    n_acc = out[2].n_acc
    # ---- per-stage device times (events between the kernels: serialises the programmatically dependent launches, so the
    #      stages add up to more than a step of region 1; used for the roofline of the dominant kernel only) -------------
    m.set_timing(1)
    stage = {}
    for _ in range(args.steps):
        flush_l2()
        step(dev_ptr)
        for k_, v in m.timing().items():
            stage[k_] = stage.get(k_, 0.0) + v
    # ---- timed region 2: end to end from pinned host memory through the C ABI (wall clock of the blocking call) -------
    m.set_timing(0)
    gc_off = not os.environ.get("MINS_BENCH_KEEP_GC")
    if gc_off:
        # a collection inside a ~160 us wall-clock region is a host artefact, not the call's cost; collect BEFORE the warm-up
        # steps of this mode (a full collection leaves the host caches cold: the step right after it measured 35 us slower)
        gc.collect()
        gc.disable()
    for _ in range(3):  # warm-up of this mode (host input, no event timers)
        flush_l2()
        step(host_ptr)
    barrier()
    e2e_s, e2e_steps = 0.0, []
    for _ in range(args.steps):
        flush_l2()
        t0 = time.perf_counter()
        step(host_ptr)
        dt_ = time.perf_counter() - t0
        e2e_s += dt_
        e2e_steps.append(dt_)
        sampler.sample()
    # ---- the same call for the other kinds of input an integrator may hand over (median of `steps` calls each; informational,
    #      in e2e.by_input): pageable packed xyz (stride 3: one host memcpy into the handle's pinned staging buffer, then the
    #      kernel reads that) and device-resident input as wall clock ------------------------------------------------------------
    by_input = {"pinned_stride4_ms": float(np.median(e2e_steps) * 1e3)}
    xyz3 = np.ascontiguousarray(w["raw"][:, :3])

    def step3():
        rc = lib.ml_scan_update(h, xyz3.ctypes.data, 3, n, w["t_rel"], op.min_range, op.max_range, 0.0, T.ctypes.data, C.byref(cns),
                                P.ctypes.data, out[0].ctypes.data, out[1].ctypes.data, C.byref(out[2]))
        if rc != 0:
            raise RuntimeError(f"ml_scan_update -> {rc}: {lib.ml_last_error(h).decode()}")

    for name, fn in (("pageable_stride3_ms", step3), ("device_resident_ms", lambda: step(dev_ptr))):
        ts_ = []
        for it in range(3 + args.steps):
            flush_l2()
            t0 = time.perf_counter()
            fn()
            if it >= 3:
                ts_.append(time.perf_counter() - t0)
        by_input[name] = float(np.median(ts_) * 1e3)
    by_input["note"] = ("wall clock of the blocking ml_scan_update call; the C++ shim (mins_gpu::UpdaterLidar) queues scans in pooled "
                        "pinned PointXYZ buffers (ml_host_alloc), so its update() call is the pinned_stride4 case")
    if gc_off:
        gc.enable()
    barrier()
    clocks = sampler.finish()

    tt = torch.tensor([dev_ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    pts = torch.tensor([float(n * args.steps)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(pts, op=dist.ReduceOp.SUM)
    dev_ms_max, e2e_ms_max = float(tt[0]), float(tt[1])
    by_rank = [[e2e_s * 1e3 / args.steps, float(np.max(e2e_steps) * 1e3), float(np.argmax(e2e_steps))]]
    if world > 1:
        mine = torch.tensor(by_rank[0], dtype=torch.float64, device="cuda")
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        by_rank = [[float(v) for v in x] for x in allr]
    total_pts = float(pts[0])
    if rank == 0:
        value = total_pts / (dev_ms_max * 1e-3)
        e2e = total_pts / (e2e_ms_max * 1e-3)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
        fused = stage.get("upload_filter", 0.0) == 0.0 and stage.get("bin", 0.0) == 0.0
        one_kernel = fused and stage.get("rows_reduce", 0.0) == 0.0  # tile.cuh: the whole update is one launch, timed under "knn"
        n_parts = min(-(-n // 448), 296) if one_kernel else 148
        kern = {"knn": ("k_scan_tile<5,3> (range filter + transform + exact kNN + plane fit + residual/Jacobian + chi2 + reduction: the whole "
                        "scan update, one kernel)" if one_kernel else
                        "k_front_knn<5> (range filter + transform + cell grouping + exact kNN, one persistent kernel)") if fused else "k_knn<5>",
                "rows_reduce": "k_rows_reduce<5,3> (plane fit + residual/Jacobian + chi2 + reduction)",
                "bin": "k_unit_alloc + k_query_scatter", "upload_filter": "k_range_filter_bin (range filter + transform + query binning)"}
        cand = {k_: stage.get(k_, 0.0) / args.steps for k_ in kern}
        dom = max(cand, key=cand.get)
        ach = ALG_BYTES_PER_POINT * n / (cand[dom] * 1e-3) / 1e9
        traffic = None
        try:
            prof = json.load(open(os.path.join(ROOT, "profiles", "ncu_summary.json")))
            key = "scan_tile" if (one_kernel and dom == "knn") else {"upload_filter": "front"}.get(dom, dom)
            traffic = prof.get(key, {}).get("dram_bytes_per_launch")
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "warmup_steps_run": n_w,
            "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 kNN / f64 rows", "data": "synthetic",
            "config": workload_desc(w, world, n, n_acc),
            "clocks": clocks,
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(n * 16 + C.sizeof(capi.ml_scan_consts) + hx * hx * 8 + 32),
                    "d2h_bytes_per_step": int(n_parts * (8 * 57 + 32)), "ms_per_step": e2e_ms_max / args.steps,
                    "step_ms_min_median_max": [float(np.min(e2e_steps) * 1e3), float(np.median(e2e_steps) * 1e3), float(np.max(e2e_steps) * 1e3)],
                    "ms_mean_max_argmax_by_rank": by_rank, "host_cpu_affinity_rank0": affinity, "by_input": by_input,
                    "h2d": "none as a copy: the kernel reads the pinned scan over PCIe itself (TMA bulk copies, one 7 KB tile per CTA)",
                    "d2h": "none as a copy: every CTA writes its reduced system and a stamp into pinned host memory over PCIe; the "
                           "library's host tail waits for the stamps, adds the systems in CTA order and applies the chain rule T^T G T"},
            "gpu_launches": int(launches),
            "stage_ms_per_step": {k_: v / args.steps for k_, v in stage.items()},
            "wall_ms_total_incl_l2_flush": wall_total * 1e3,
            "roofline": {"bound": "hbm", "kernel": kern[dom], "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": traffic, "peak_source": peak_src,
                         "note": "algorithmic 96 B/point x points per launch / that kernel's CUDA-event time; the map (16 MB) is "
                                 "L2-resident by nature, the path is latency/issue bound, not HBM bound (DESIGN.md)"},
        }
        if world == 1 and not args.no_aux:
            try:
                line["aux"] = aux_measurements(m, w, op, step, host_ptr, out, flush_l2, hx)
            except Exception as e:  # informational block, never required for the headline
                line["aux"] = {"error": repr(e)}
            try:
                line["aux"]["config1_2xVLP16_200k_map"] = aux_config1(local_rank)
            except Exception as e:
                line["aux"]["config1_2xVLP16_200k_map"] = {"error": repr(e)}
            try:
                line["aux"]["monte_carlo_cpp_harness"] = aux_monte_carlo()
            except Exception as e:
                line["aux"]["monte_carlo_cpp_harness"] = {"error": repr(e)}
        if world == 1 and not args.no_cpu_baseline:
            try:
                val, ms, _, sample, n_acc_cpu = cpu_reference_run(w, 3, 1, threads)
                line["cpu_baseline"] = {"value": val, "unit": UNIT, "cores": threads, "kind": "reference", "sample": sample,
                                        "ms_per_step": ms, "accepted_points_per_step": n_acc_cpu}
            except Exception as e:  # the baseline is reported, never required for the GPU number
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": threads, "kind": "reference", "sample": f"unavailable: {e}"}
        emit(line)
    m.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
