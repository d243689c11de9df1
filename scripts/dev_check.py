"""GPU-box diagnostics (not a test): parity of every stage against the oracle with mismatch statistics, and timings."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from mins_b200 import synth
from mins_b200.lidar import LidarMap, OptionsLidar, info_to_sqrt
from oracle import pyoracle as po


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def check(shape, variant="plain", k=5, brute_n=4000, **kw):
    t0 = time.time()
    op = OptionsLidar(raw_do_downsample=False, map_ngbr_num=k)
    w = synth.make_workload(shape, op=op, variant=variant, **kw)
    print(f"[{shape}/{variant}/k{k}] gen {time.time()-t0:.1f}s rays={len(w['raw'])} map={len(w['map'])} hx={w['total_hx']}", flush=True)
    m = LidarMap(max_scan_points=max(len(w['raw']), 1024), max_map_points=len(w['map']) + len(w['raw']) + 1024)
    m.Build(w['map'])
    rc, HtH, Htr, cnt = m.scan_update(w['raw'], w['t_rel'], op, w['T_LtoM'], w['consts'], w['P'])
    print("  rc", rc, cnt.as_dict(), "timing", {k_: round(v, 4) for k_, v in m.timing().items()}, flush=True)
    sc = po.range_filter(w['raw'], w['t_rel'], op.max_range, op.min_range)
    g_sc = m.scan(0)
    print("  range filter equal:", np.array_equal(sc, g_sc), len(sc), len(g_sc))
    sm = po.transform_f32(sc, w['T_LtoM'])
    g_sm = m.scan(1)
    print("  transform equal:", np.array_equal(sm, g_sm))
    idx, d2, ok = m.get_neighbors(k, op.map_ngbr_max_d)
    # brute force on a subset of queries
    rng = np.random.default_rng(0)
    sub = np.sort(rng.choice(len(sm), min(brute_n, len(sm)), replace=False))
    t0 = time.time()
    oidx, od2, ook = po.knn_brute(w['map'], sm[sub], k, op.map_ngbr_max_d)
    print(f"  brute {time.time()-t0:.1f}s ok-mask equal: {np.array_equal(ok[sub], ook)} idx equal: {np.array_equal(idx[sub][ook], oidx[ook])} "
          f"d2 equal: {np.array_equal(d2[sub][ook], od2[ook])} accept frac {ook.mean():.3f}")
    if not np.array_equal(idx[sub][ook], oidx[ook]):
        bad = np.where((idx[sub] != oidx).any(1) & ook)[0]
        print("   mismatching queries:", len(bad), "e.g.", idx[sub][bad[0]], oidx[bad[0]], d2[sub][bad[0]], od2[bad[0]])
    # linsys with the GPU's own kNN indices on the full scan (oracle rows in C++, 8 threads)
    t0 = time.time()
    ref = po.scan_linsys(w['consts'], w['P'], w['map'], sc, sm, knn_idx=idx)
    print(f"  oracle linsys {time.time()-t0:.1f}s cnt {ref['cnt'].as_dict()}")
    print("  HtH rel", rel(HtH, ref['HtH']), "Htr rel", rel(Htr, ref['Htr']))
    rows, h = m.point_rows()
    orow = ref['rows']
    print("  status equal:", np.array_equal(rows['status'], orow['status']), np.bincount(rows['status'], minlength=5), np.bincount(orow['status'], minlength=5))
    both = (rows['status'] == 0) & (orow['status'] == 0)
    if both.any():
        pe = np.abs(rows['plane'][both] - orow['plane'][both]).max(1) / np.abs(orow['plane'][both]).max(1)
        print("  plane rel max", pe.max(), "res rel max", (np.abs(rows['res'][both] - orow['res'][both]) / np.maximum(np.abs(orow['res'][both]), 1e-3)).max(),
              "chi2 rel max", (np.abs(rows['chi2'][both] - orow['chi2'][both]) / np.maximum(np.abs(orow['chi2'][both]), 1e-6)).max())
        print("  w rel max", (np.abs(rows['w'][both] - orow['w'][both]) / orow['w'][both]).max(), "vr abs-rel", (np.abs(rows['vr'][both] - orow['vr'][both]) / np.maximum(np.abs(orow['vr'][both]), 1e-6)).max(),
              "h rel", rel(h[both], ref['h'][both]))
    Rc, rcv = info_to_sqrt(HtH, Htr)
    dx1, P1 = po.ekf_update(w['P'], Rc, rcv)
    ref2 = po.scan_linsys(w['consts'], w['P'], w['map'], sc, sm, knn_idx=idx, want_compress=True)
    dx0, P0 = po.ekf_update(w['P'], ref2['Hc'], ref2['rc'])
    print("  EKF dx rel", rel(dx1, dx0), "dP rel", rel(P1 - w['P'], P0 - w['P']), "rank", len(rcv))
    return m, w


def timing(shape, reps=20, **kw):
    op = OptionsLidar(raw_do_downsample=False)
    w = synth.make_workload(shape, op=op, **kw)
    m = LidarMap(max_scan_points=len(w['raw']), max_map_points=len(w['map']) + 2 * len(w['raw']))
    t0 = time.time(); m.Build(w['map']); print(f"[{shape}] map build wall {1e3*(time.time()-t0):.2f} ms")
    hx = w['total_hx']
    out = (np.zeros((hx, hx)), np.zeros(hx), __import__('mins_b200.capi', fromlist=['x']).ml_counters())
    for timing_on in (1, 0):
        m.set_timing(timing_on)
        ts = []
        for r in range(reps):
            t0 = time.perf_counter()
            m.scan_update(w['raw'], w['t_rel'], op, w['T_LtoM'], w['consts'], w['P'], out=out)
            ts.append(time.perf_counter() - t0)
        ts = np.array(ts[3:])
        print(f"  timing_on={timing_on} e2e wall median {1e3*np.median(ts):.3f} ms min {1e3*ts.min():.3f} -> {len(w['raw'])/np.median(ts)/1e6:.1f} Mpts/s; stages {m.timing()}", flush=True)
    t0 = time.perf_counter(); n = m.register_scan(op); t1 = time.perf_counter()
    print(f"  register_scan wall {1e3*(t1-t0):.2f} ms -> map {n}")
    return m, w


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "small"):
        check("G-small")
        check("G-small", variant="calib_dt", k=5)
        check("G-small", variant="overlap", k=7)
        check("G-small", variant="at_clone", k=3)
    if which in ("all", "vlp"):
        check("G-vlp", brute_n=3000)
    if which in ("all", "time"):
        timing("G-vlp")
        timing("G-64")
