"""GPU parity tests proper: every stage of the CUDA path, called through the C ABI (mins_b200.lidar -> ctypes ->
libmins_lidar.so), against the CPU oracle on the same seeded inputs and against the committed golden fixtures.

Bars: bit-exact for the fp32/integer stages (range filter, transform, kNN indices AND distances, map
insert point set, re-anchor, deskew); plane / residual / Jacobian rows and the reduced system within 1e-9 relative
(fp64; the reduction order differs); voxel-grid centroids within 4 ulp*sqrt(n) with identical voxel membership and order
(PCL's own sort is unstable, so bit-equality is not defined even for the reference).
"""
import numpy as np
import pytest

from conftest import check_voxel_contract, load_golden, rel

pytestmark = pytest.mark.gpu

TOL = 1e-9


def as_xyzi(xyz, w=None):
    a = np.zeros((len(xyz), 4), np.float32)
    a[:, :3] = xyz
    if w is not None:
        a[:, 3] = w
    return a


def sort_rows(a):
    return a[np.lexsort((a[:, 3], a[:, 2], a[:, 1], a[:, 0]))]


@pytest.fixture()
def make_map(cuda_ok):
    from mins_b200.lidar import LidarMap
    made = []

    def mk(**kw):
        kw.setdefault("max_scan_points", 1 << 16)
        kw.setdefault("max_map_points", 1 << 18)
        m = LidarMap(**kw)
        made.append(m)
        return m

    yield mk
    for m in made:
        m.close()


# ------------------------------------------------------------------------------------------------ kNN
@pytest.mark.parametrize("name", ["nabo_scan3d.npz", "nabo_scan3d_large.npz"])
@pytest.mark.parametrize("k", [1, 5, 10])
def test_knn_bit_exact_on_libnabo_fixtures(make_map, po, name, k):
    g = load_golden(name)
    cloud, q = as_xyzi(g["cloud"]), as_xyzi(g["queries"])
    m = make_map()
    m.Build(cloud)
    m.set_scan(q)
    m.transform_to_map(np.eye(4, dtype=np.float32))
    # bounded radius as get_neighbors uses it; the fixture's k-th distances straddle it
    for max_d2 in (np.float32(1.0), np.float32(0.04), np.float32(25.0)):
        idx, d2, ok = m.get_neighbors(k, float(max_d2))
        oidx, od2, ook = po.knn_brute(cloud, q, k, max_d2)
        assert np.array_equal(ok, ook)
        assert np.array_equal(idx[ook], oidx[ook])
        assert np.array_equal(d2[ook], od2[ook])
        # and against the reference tree's own distances
        assert np.array_equal(d2[ook], g["ref_d2"][ook][:, :k])


@pytest.mark.parametrize("cell", [0.25, 0.5, 1.0, 3.0])
def test_knn_independent_of_grid_cell(make_map, po, workload, cell):
    """the grid is an acceleration structure only: results must not depend on its resolution"""
    w = workload("G-small")
    op = w["op"]
    m = make_map(grid_cell=cell)
    m.Build(w["map"])
    m.feed_scan(w["raw"], w["t_rel"], op)
    m.transform_to_map(w["T_LtoM"])
    idx, d2, ok = m.get_neighbors(5, op.map_ngbr_max_d)
    sm = m.scan(1)
    oidx, od2, ook = po.knn_brute(w["map"], sm, 5, np.float32(op.map_ngbr_max_d))
    assert np.array_equal(ok, ook) and np.array_equal(idx[ook], oidx[ook]) and np.array_equal(d2[ook], od2[ook])


def test_knn_tie_break_is_ascending_map_index(make_map, po):
    """duplicate map points and lattice points at exactly equal distances: (d2, index) ascending — libnabo brute-force
    semantics (brute_force_cpu.cpp:86-93), the stated rule"""
    g = np.stack(np.meshgrid(*[np.arange(-3, 4)] * 3, indexing="ij"), -1).reshape(-1, 3).astype(np.float32) * 0.25
    cloud = as_xyzi(np.concatenate([g, g[:50]]))  # 50 exact duplicates with larger indices
    rng = np.random.default_rng(0)
    q = as_xyzi(np.concatenate([g[rng.choice(len(g), 64)], np.zeros((1, 3), np.float32)]))  # on lattice nodes: many ties
    m = make_map()
    m.Build(cloud)
    m.set_scan(q)
    m.transform_to_map(np.eye(4, dtype=np.float32))
    for k in (1, 5, 7, 10):
        idx, d2, ok = m.get_neighbors(k, 4.0)
        oidx, od2, ook = po.knn_brute(cloud, q, k, np.float32(4.0))
        assert ook.all() and np.array_equal(ok, ook)
        assert np.array_equal(idx, oidx) and np.array_equal(d2, od2)


def test_knn_edge_cases(make_map, po):
    rng = np.random.default_rng(1)
    cloud = as_xyzi(rng.uniform(-5, 5, (4, 3)))
    q = as_xyzi(rng.uniform(-5, 5, (100, 3)))
    m = make_map()
    # no map yet: soft status, like tree->initialized() == false
    m.set_scan(q)
    m.transform_to_map(np.eye(4, dtype=np.float32))
    from mins_b200 import capi
    rc = m.lib.ml_knn(m.h, 5, 1.0, None, None, None, None)
    assert rc == capi.ML_SOFT_NO_MAP
    # fewer map points than k: nothing accepted, partial lists padded with -1
    m.Build(cloud)
    idx, d2, ok = m.get_neighbors(5, 1e6)
    oidx, od2, ook = po.knn_brute(cloud, q, 5, np.float32(1e6))
    assert not ok.any() and np.array_equal(idx, oidx)
    # queries far outside the map's bounding box
    far = as_xyzi(rng.uniform(500, 600, (64, 3)))
    m.set_scan(far)
    m.transform_to_map(np.eye(4, dtype=np.float32))
    idx, d2, ok = m.get_neighbors(3, 1.0)
    assert not ok.any() and (idx == -1).all()
    # empty scan
    m.set_scan(np.zeros((0, 4), np.float32))
    m.transform_to_map(np.eye(4, dtype=np.float32))
    idx, d2, ok = m.get_neighbors(3, 1.0)
    assert len(idx) == 0
    # k out of range
    assert m.lib.ml_knn(m.h, 11, 1.0, None, None, None, None) == capi.ML_ERR_ARG
    assert m.lib.ml_knn(m.h, 0, 1.0, None, None, None, None) == capi.ML_ERR_ARG


# ----------------------------------------------------------------------------------- scan pre-processing
def test_range_filter_and_transform_bit_exact(make_map, po, workload):
    w = workload("G-vlp")
    op = w["op"]
    m = make_map()
    for stride in (4, 3):
        raw = w["raw"] if stride == 4 else np.ascontiguousarray(w["raw"][:, :3])
        n = m.feed_scan(raw, w["t_rel"], op)
        sc = po.range_filter(w["raw"], w["t_rel"], op.max_range, op.min_range)
        assert n == len(sc) and 0 < n < len(raw)  # rays without a return are (0,0,0): filtered by min_range
        assert np.array_equal(m.scan(0), sc) and np.array_equal(m.scan(2), sc)
        m.transform_to_map(w["T_LtoM"])
        assert np.array_equal(m.scan(1), po.transform_f32(sc, w["T_LtoM"]))


def test_range_filter_edge_cases(make_map, po):
    from mins_b200.lidar import OptionsLidar
    op = OptionsLidar(min_range=1.0, max_range=10.0)
    m = make_map()
    # exactly on the bounds is kept (`n > max || n < min` rejects, LidarTypes.cpp:39); all-rejected and empty scans
    pts = np.array([[1, 0, 0], [10, 0, 0], [0.999, 0, 0], [10.001, 0, 0], [0, 0, 0], [3, 4, 0]], np.float32)
    assert m.feed_scan(pts, 0.5, op) == 3
    assert np.array_equal(m.scan(0), po.range_filter(pts, 0.5, 10.0, 1.0))
    assert m.feed_scan(np.zeros((100, 3), np.float32), 0.5, op) == 0
    assert len(m.scan(0)) == 0
    assert m.feed_scan(np.zeros((0, 3), np.float32), 0.5, op) == 0


@pytest.mark.parametrize("leaf", [0.3, 1.0, 2.0])
def test_voxel_downsample_membership_order_and_centroids(make_map, po, workload, leaf):
    """pcl::VoxelGrid semantics (LidarHelper.cpp:89-94): one centroid per occupied voxel, ascending voxel index."""
    w = workload("G-vlp")
    op = w["op"]
    m = make_map()
    m.feed_scan(w["raw"], w["t_rel"], op)
    sc = m.scan(0)
    n_ds = m.downsample(leaf)
    out = m.scan(0)
    ref, vid = po.voxel_downsample(sc, leaf)
    assert n_ds == len(ref) == len(out)
    assert (np.diff(vid.astype(np.int64)) > 0).all()
    # membership/order: every GPU centroid lies in the voxel the oracle assigns to that output slot
    inv = np.float32(1.0) / np.float32(leaf)
    mn = np.floor(sc[:, :3].min(0) * inv).astype(np.int64)
    dv = np.floor(sc[:, :3].max(0) * inv).astype(np.int64) - mn + 1
    # centroid tolerance: 4 ulp * sqrt(n) of the coordinate magnitude
    counts = np.bincount(np.searchsorted(vid, ((np.floor(sc[:, 0] * inv).astype(np.int64) - mn[0])
                                               + (np.floor(sc[:, 1] * inv).astype(np.int64) - mn[1]) * dv[0]
                                               + (np.floor(sc[:, 2] * inv).astype(np.int64) - mn[2]) * dv[0] * dv[1])),
                         minlength=len(vid))
    assert counts.sum() == len(sc) and (counts > 0).all()
    tol = 4 * np.finfo(np.float32).eps * np.sqrt(counts)[:, None] * np.maximum(np.abs(ref[:, :3]).max(1), 1.0)[:, None]
    assert (np.abs(out[:, :3] - ref[:, :3]) <= tol).all()
    assert np.array_equal(out[:, 3], ref[:, 3])  # all points share one time stamp: the mean is exact


def test_voxel_downsample_single_point_voxels_are_exact(make_map, po):
    rng = np.random.default_rng(2)
    pts = as_xyzi(rng.uniform(-50, 50, (5000, 3)), 1.5)
    m = make_map()
    m.set_scan(pts)
    n = m.downsample(0.01)  # (almost) every point alone in its voxel -> output = input sorted by voxel index
    ref, _ = po.voxel_downsample(pts, 0.01)
    assert n == len(ref)
    assert np.array_equal(m.scan(0), ref)


def test_deskew_per_point_transforms_bit_exact(make_map, po):
    rng = np.random.default_rng(4)
    n = 3000
    pts = as_xyzi(rng.uniform(-30, 30, (n, 3)), 0.25)
    from mins_b200.synth import exp_so3
    T = np.zeros((n, 3, 4), np.float32)
    for i in range(n):
        a = i / n
        T[i, :, :3] = exp_so3(np.array([0.01, -0.02, 0.05]) * a)
        T[i, :, 3] = np.array([0.5, 0.02, -0.01]) * a
    m = make_map()
    m.set_scan(pts)
    m.deskew(T.reshape(n, 12), n)
    assert np.array_equal(m.scan(0), po.deskew(pts, T.reshape(n, 12)))


# -------------------------------------------------------------------------------------- rows and reduction
@pytest.mark.parametrize("shape,variant,k", [("G-small", "plain", 5), ("G-small", "calib", 5), ("G-small", "calib_dt", 5),
                                             ("G-small", "overlap", 7), ("G-small", "at_clone", 3), ("G-small", "plain", 10),
                                             ("G-vlp", "plain", 5)])
def test_linsys_rows_and_reduction_match_oracle(make_map, po, workload, shape, variant, k):
    from mins_b200.lidar import info_to_sqrt
    w = workload(shape, variant=variant, k=k)
    op = w["op"]
    m = make_map(max_map_points=1 << 19)
    m.Build(w["map"])
    rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert rc == 0
    sc, sm = m.scan(0), m.scan(1)
    ref = po.scan_linsys(w["consts"], w["P"], w["map"], sc, sm, want_compress=True)  # oracle does its own brute-force kNN
    # decisions: identical status per point (no point sits within 1e-9 of a gate on these fixtures)
    rows, h = m.point_rows()
    assert np.array_equal(rows["status"], ref["rows"]["status"])
    gc, oc = cnt.as_dict(), ref["cnt"].as_dict()
    for key in ("n_points", "n_knn_fail", "n_plane_fail", "n_badmatch", "n_chi_rej", "n_acc"):
        assert gc[key] == oc[key]
    assert gc["n_acc"] > 0.4 * gc["n_points"]
    assert gc["sum_chi"] == pytest.approx(oc["sum_chi"], rel=TOL) and gc["sum_res2"] == pytest.approx(oc["sum_res2"], rel=TOL)
    # per-point values
    both = rows["status"] == 0
    o = ref["rows"]
    assert (np.abs(rows["plane"][both] - o["plane"][both]).max(1) <= TOL * np.abs(o["plane"][both]).max(1)).all()
    assert np.abs(rows["res"][both] - o["res"][both]).max() <= TOL * max(np.abs(o["res"][both]).max(), 1.0)
    assert np.abs(rows["chi2"][both] - o["chi2"][both]).max() <= TOL * max(np.abs(o["chi2"][both]).max(), 1.0)
    assert np.abs(rows["raw_n"][both] - o["raw_n"][both]).max() <= TOL * np.abs(o["raw_n"][both]).max()
    assert np.abs(rows["w"][both] - o["w"][both]).max() <= TOL
    assert rel(h[both], ref["h"][both]) <= TOL
    # reduced system
    assert rel(HtH, ref["HtH"]) <= TOL and rel(Htr, ref["Htr"]) <= TOL
    assert np.array_equal(HtH, HtH.T)
    # and the EKF update it produces equals the one from the reference's Givens-compressed stack
    Rc, rcv = info_to_sqrt(HtH, Htr)
    dx1, P1 = po.ekf_update(w["P"], Rc, rcv)
    dx0, P0 = po.ekf_update(w["P"], ref["Hc"], ref["rc"])
    assert rel(dx1, dx0) <= 1e-8 and rel(P1 - w["P"], P0 - w["P"]) <= 1e-8


def test_linsys_with_downsampled_scan(make_map, po, workload):
    w = workload("G-vlp", downsample=True)
    op = w["op"]
    m = make_map()
    m.Build(w["map"])
    rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert rc == 0 and 0 < cnt.n_points < 0.5 * w["n_rays"]
    sc, sm = m.scan(0), m.scan(1)  # identical query inputs for both sides (SURVEY.md §7 "downsample is not bit-reproducible")
    assert np.array_equal(sm, po.transform_f32(sc, w["T_LtoM"]))
    ref = po.scan_linsys(w["consts"], w["P"], w["map"], sc, sm)
    assert cnt.n_acc == ref["cnt"].n_acc
    assert rel(HtH, ref["HtH"]) <= TOL and rel(Htr, ref["Htr"]) <= TOL


def test_linsys_soft_and_error_statuses(make_map, workload):
    from mins_b200 import capi
    from mins_b200.lidar import MLError
    w = workload("G-small")
    op = w["op"]
    m = make_map()
    # map not initialised -> soft status 2 (caller keeps the scan, UpdaterLidar.cpp:91)
    rc, *_ = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert rc == capi.ML_SOFT_NO_MAP
    m.Build(w["map"])
    # a scan that matches nothing -> "no rows" soft status 1 (UpdaterLidar::update returns false, UpdaterLidar.cpp:398)
    far = w["raw"].copy()
    far[:, :3] = far[:, :3] * 0 + np.array([50.0, 50.0, 60.0], np.float32)
    rc, HtH, Htr, cnt = m.scan_update(far, w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert rc == capi.ML_SOFT_NO_ROWS and cnt.n_acc == 0 and not HtH.any() and not Htr.any()
    # empty scan
    rc, HtH, Htr, cnt = m.scan_update(np.zeros((0, 4), np.float32), w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert rc == capi.ML_SOFT_NO_ROWS and cnt.n_points == 0
    # capacity exceeded -> fatal
    small = make_map(max_scan_points=100, max_map_points=1000)
    with pytest.raises(MLError) as e:
        small.feed_scan(w["raw"], 0.0, op)
    assert e.value.code == capi.ML_ERR_CAPACITY
    with pytest.raises(MLError):
        small.Build(w["map"])
    # bad constants
    import copy
    c2 = copy.copy(w["consts"])
    c2.total_hx = 65
    with pytest.raises(MLError) as e:
        m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], c2, np.zeros((65, 65)))
    assert e.value.code == capi.ML_ERR_ARG


def test_chi2_gate_rejects_like_the_reference(make_map, po, workload):
    """tighten chi2_mult so a sizeable fraction is rejected; decisions must still agree point by point"""
    import copy
    w = workload("G-small")
    c2 = copy.copy(w["consts"])
    c2.chi2_mult = 0.02 / 7.814727903251179
    m = make_map()
    m.Build(w["map"])
    rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], w["op"], w["T_LtoM"], c2, w["P"])
    ref = po.scan_linsys(c2, w["P"], w["map"], m.scan(0), m.scan(1))
    rows, _ = m.point_rows(want_h=False)
    assert cnt.n_chi_rej > 50 and cnt.n_acc > 50
    assert np.array_equal(rows["status"], ref["rows"]["status"])
    assert rel(HtH, ref["HtH"]) <= TOL


def test_linsys_is_deterministic_and_additive(make_map, workload):
    """size-independent properties: bit-identical repeats; linsys(A u B) = linsys(A) + linsys(B)"""
    w = workload("G-vlp")
    op = w["op"]
    m = make_map()
    m.Build(w["map"])
    r1 = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    r2 = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert np.array_equal(r1[1], r2[1]) and np.array_equal(r1[2], r2[2])
    half = len(w["raw"]) // 2
    ra = m.scan_update(w["raw"][:half], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    rb = m.scan_update(w["raw"][half:], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert ra[3].n_acc + rb[3].n_acc == r1[3].n_acc
    assert rel(ra[1] + rb[1], r1[1]) <= 1e-12 and rel(ra[2] + rb[2], r1[2]) <= 1e-11


# ---------------------------------------------------------------------------------------------- map
def test_map_add_equals_reference_add_points_golden(make_map, po):
    g = load_golden("ikd_add_points.npz")
    m = make_map()
    m.Build(g["base"])
    assert np.array_equal(m.flatten(), g["base"])  # Build keeps every point, in order
    n = m.Add_Points(g["add"], True, float(g["ds"]))
    out = m.flatten()
    assert n == len(out) == len(g["result_sorted"])
    assert np.array_equal(sort_rows(out), g["result_sorted"])            # the reference's own result set
    assert np.array_equal(out, po.map_insert(g["base"], g["add"], True, float(g["ds"])))  # and the flat-list order


def test_register_scan_equals_oracle_and_knn_sees_the_new_map(make_map, po, workload):
    w = workload("G-vlp")
    op = w["op"]
    m = make_map(max_map_points=1 << 19)
    m.Build(w["map"])
    m.feed_scan(w["raw"], w["t_rel"], op)
    m.transform_to_map(w["T_LtoM"])
    sm = m.scan(1)
    n = m.register_scan(op)
    ref = po.map_insert(w["map"], sm, True, op.map_downsample_size)
    out = m.flatten()
    assert n == len(ref) and np.array_equal(out, ref)
    # the kNN grid was rebuilt over the new map
    idx, d2, ok = m.get_neighbors(5, op.map_ngbr_max_d)
    sub = np.arange(0, len(sm), 37)
    oidx, od2, ook = po.knn_brute(ref, sm[sub], 5, np.float32(op.map_ngbr_max_d))
    assert np.array_equal(ok[sub], ook) and np.array_equal(idx[sub][ook], oidx[ook]) and np.array_equal(d2[sub][ook], od2[ook])
    # idempotence: registering the same scan again changes nothing (every voxel's winner is already in the map)
    n2 = m.register_scan(op)
    assert n2 == n and np.array_equal(sort_rows(m.flatten()), sort_rows(ref))


def test_map_add_without_downsample_and_capacity(make_map, po):
    from mins_b200.lidar import MLError
    rng = np.random.default_rng(6)
    a, b = as_xyzi(rng.uniform(-5, 5, (500, 3)), 0.1), as_xyzi(rng.uniform(-5, 5, (300, 3)), 0.2)
    m = make_map(max_scan_points=400, max_map_points=700)
    m.Build(a)
    assert m.Add_Points(b[:100], False, 0.3) == 600
    assert np.array_equal(m.flatten(), np.concatenate([a, b[:100]]))
    with pytest.raises(MLError):
        m.Add_Points(b[:200], False, 0.3)  # 800 > max_map_points
    m.reset()
    assert not m.initialized() and m.size() == 0


def test_deferred_map_update_equals_the_waiting_one(make_map, po, workload):
    """n_live == NULL (what the C++ shim passes: register_scan returns void): the merge and the grid rebuild are only queued;
    the next calls see the same map as after the waiting variant, bit for bit, including a scan update issued right behind."""
    import ctypes as C
    from mins_b200.lidar import OptionsLidar
    w = workload("G-small")
    op = OptionsLidar(map_do_downsample=True, map_downsample_size=0.3)
    T = np.eye(4)
    T[:3, 3] = [0.05, -0.02, 0.01]
    res = []
    for wait in (True, False):
        m = make_map()
        m.Build(w["map"])
        m.feed_scan(w["raw"], w["t_rel"], op)
        m.transform_to_map(w["T_LtoM"])
        n = m.register_scan(op, wait=wait)
        assert (n is None) == (not wait)
        # straight behind the (possibly still running) merge: a full scan update against the new map
        rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
        size_after_insert = m.size()
        n2 = m.propagate_map_to_newest_clone(T, 1.0, op, wait=wait)
        res.append((size_after_insert, HtH.copy(), Htr.copy(), cnt.as_dict(), m.flatten(), m.size()))
    a, b = res
    assert a[0] == b[0] and a[5] == b[5] and a[3] == b[3]
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2]) and np.array_equal(a[4], b[4])


def test_deferred_map_update_reports_capacity_on_the_next_call(make_map):
    from mins_b200 import capi
    from mins_b200.lidar import MLError
    import ctypes as C
    rng = np.random.default_rng(8)
    a, b = as_xyzi(rng.uniform(-5, 5, (500, 3)), 0.1), as_xyzi(rng.uniform(-5, 5, (300, 3)), 0.2)
    m = make_map(max_scan_points=400, max_map_points=700)
    m.Build(a)
    bb = np.ascontiguousarray(b, np.float32)
    assert m.lib.ml_map_add(m.h, bb.ctypes.data, 300, 0, 0.3, None) == capi.ML_OK       # 800 > 700, not known yet
    with pytest.raises(MLError) as e:
        m.size()
    assert e.value.code == capi.ML_ERR_CAPACITY
    assert m.size() == 500 and np.array_equal(m.flatten(), a)                             # the map is unchanged and usable
    assert m.Add_Points(b[:100], False, 0.3) == 600


@pytest.mark.parametrize("do_ds", [False, True])
def test_reanchor_equals_oracle(make_map, po, workload, do_ds):
    """LidarHelper::propagate_map_to_newest_clone (LidarHelper.cpp:209-266)"""
    from mins_b200.lidar import OptionsLidar
    from mins_b200.synth import exp_so3
    w = workload("G-small")
    op = OptionsLidar(map_do_downsample=do_ds, map_downsample_size=0.6, map_decay_time=3.0, map_decay_dist=20.0)
    T = np.eye(4)
    T[:3, :3] = exp_so3(np.array([0.01, 0.02, -0.05]))
    T[:3, 3] = [-1.0, 0.05, 0.02]
    m = make_map()
    m.Build(w["map"])
    n = m.propagate_map_to_newest_clone(T, 5.5, op)
    ref = po.reanchor(w["map"], T, 5.5, op.map_decay_time, op.map_decay_dist, do_ds, op.map_downsample_size)
    out = m.flatten()
    assert n == len(ref) and 0 < n < len(w["map"])
    if not do_ds:
        assert np.array_equal(out, ref)
    else:  # centroids: fp32 accumulation order inside a voxel is not defined by PCL either -> the stated voxel contract
        # (identical membership and order, 4 ulp sqrt(n)), carried through the decay filters and the fp64 transform
        def post(c):
            keep = ~(c[:, 3].astype(np.float64) + op.map_decay_time < 5.5)
            nrm = np.sqrt((c[:, 0] * c[:, 0] + (c[:, 1] * c[:, 1] + c[:, 2] * c[:, 2]).astype(np.float32)).astype(np.float32))
            keep &= ~(nrm > op.map_decay_dist)
            q = c[keep].copy()
            q[:, :3] = (c[keep, :3].astype(np.float64) @ T[:3, :3].T + T[:3, 3]).astype(np.float32)
            return q, keep
        check_voxel_contract(w["map"], out, op.map_downsample_size, po, post=post, slack=2.0)
    assert m.initialized()


# ------------------------------------------------------------------------------- full BASELINE sizes (properties)
def test_full_size_g64_properties(make_map, po, workload):
    """config[2]: 131 072-ray scan vs 1 M-point map. The oracle is too slow for all queries: brute force on a sample,
    size-independent properties on everything."""
    w = workload("G-64")
    op = w["op"]
    n_rays, M = w["n_rays"], len(w["map"])
    assert n_rays == 131072 and M == 1_000_000
    m = make_map(max_scan_points=n_rays, max_map_points=M + 2 * n_rays)
    m.Build(w["map"])
    rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert rc == 0
    c = cnt.as_dict()
    assert c["n_knn_fail"] + c["n_plane_fail"] + c["n_badmatch"] + c["n_chi_rej"] + c["n_acc"] == c["n_points"]
    assert c["n_acc"] > 0.4 * c["n_points"]
    sm = m.scan(1)
    idx, d2, ok = m.get_neighbors(5, op.map_ngbr_max_d)
    good = idx[ok]
    assert (np.diff(d2[ok], axis=1) >= 0).all()                      # sortedness
    assert (good >= 0).all() and (good < M).all()
    assert (np.sort(good, axis=1)[:, 1:] != np.sort(good, axis=1)[:, :-1]).all()   # no neighbour twice
    # returned distances are the true fp32 distances of the returned indices (checksum of the whole result)
    dd = w["map"][good][..., :3] - sm[ok][:, None, :3]
    d2_chk = (dd[..., 0] * dd[..., 0] + dd[..., 1] * dd[..., 1]) + dd[..., 2] * dd[..., 2]
    assert np.array_equal(d2_chk.astype(np.float32), d2[ok])
    # exactness on a sample, bit-exact against brute force
    sub = np.random.default_rng(0).choice(len(sm), 1500, replace=False)
    oidx, od2, ook = po.knn_brute(w["map"], sm[sub], 5, np.float32(op.map_ngbr_max_d))
    assert np.array_equal(ok[sub], ook) and np.array_equal(idx[sub][ook], oidx[ook]) and np.array_equal(d2[sub][ook], od2[ook])
    # reduced system against the oracle rows driven with the (verified) GPU neighbour lists
    ref = po.scan_linsys(w["consts"], w["P"], w["map"], m.scan(0), sm, knn_idx=idx)
    assert ref["cnt"].n_acc == c["n_acc"]
    assert rel(HtH, ref["HtH"]) <= TOL and rel(Htr, ref["Htr"]) <= TOL
    ev = np.linalg.eigvalsh(HtH)
    assert ev.min() >= -1e-9 * ev.max()                                # PSD
    # registration at full size: one survivor per touched voxel, map grows
    n_before = m.size()
    n_after = m.register_scan(op)
    assert n_before < n_after <= n_before + c["n_points"]


def test_one_call_downsampling_path_equals_the_separate_calls(make_map, po, workload):
    """ml_scan_update with leaf > 0 (front end that also derives the voxel grid, voxel filter whose last kernel transforms and
    bins the centroids) against ml_scan_upload -> ml_scan_downsample -> ml_scan_transform -> ml_linsys: same clouds, same
    reduced system, for pageable / pinned / device input, ragged and tiny sizes, a scan the range gate rejects entirely,
    and a leaf PCL would refuse."""
    import ctypes as C
    import torch
    from mins_b200 import capi
    w = workload("G-vlp")
    op = w["op"]
    m = make_map(max_map_points=1 << 19)
    m.Build(w["map"])
    hx = w["total_hx"]
    T = np.ascontiguousarray(w["T_LtoM"], np.float32)
    P = np.ascontiguousarray(w["P"], np.float64)
    raw = np.ascontiguousarray(w["raw"], np.float32).copy()
    raw[::3, :3] *= 1000.0
    far = raw.copy()
    far[:, :3] = 5000.0
    for leaf in (0.3, 1.0, 1e-6):
        for sub in (raw, raw[:2048 * 3 + 17], raw[:33], raw[:1], far[:4096]):
            sub = np.ascontiguousarray(sub)
            n = len(sub)
            kept = m.feed_scan(sub, w["t_rel"], op)
            n_ds = C.c_int()
            assert m.lib.ml_scan_downsample(m.h, leaf, C.byref(n_ds)) == 0
            m.transform_to_map(T)
            rc0, HtH0, Htr0, cnt0 = m.linsys(w["consts"], P)
            s0 = (m.scan(0), m.scan(1), m.scan(2))
            assert len(s0[0]) == n_ds.value <= kept
            # (centroid values against the oracle: test_voxel_downsample_membership_order_and_centroids; here only the count)
            in_range = po.range_filter(sub, w["t_rel"], op.max_range, op.min_range)
            assert len(s0[0]) == (len(po.voxel_downsample(in_range, leaf)[0]) if len(in_range) else 0)
            pinned = torch.from_numpy(sub).pin_memory()
            dev = torch.from_numpy(sub).cuda()
            for ptr in (sub.ctypes.data, pinned.data_ptr(), dev.data_ptr()):
                for _ in range(2):
                    HtH, Htr, cnt = np.zeros((hx, hx)), np.zeros(hx), capi.ml_counters()
                    rc = m.lib.ml_scan_update(m.h, ptr, 4, n, w["t_rel"], op.min_range, op.max_range, leaf, T.ctypes.data,
                                              C.byref(w["consts"]), P.ctypes.data, HtH.ctypes.data, Htr.ctypes.data, C.byref(cnt))
                    assert rc == rc0
                    assert np.array_equal(m.scan(0), s0[0]) and np.array_equal(m.scan(1), s0[1]) and np.array_equal(m.scan(2), s0[2])
                    assert cnt.as_dict() == cnt0.as_dict() and np.array_equal(HtH, HtH0) and np.array_equal(Htr, Htr0)


@pytest.mark.parametrize("path", ["auto", "one_kernel"])
def test_one_launch_front_end_equals_the_multi_call_path(make_map, po, workload, path, monkeypatch):
    """ml_scan_update's fused front end (single-pass compaction + transform + binning, zero-copy pinned input) must leave
    exactly the state the separate calls leave: same clouds, same kNN, same reduced system - for packed xyz (stride 3),
    PointXYZ (stride 4), pageable and pinned host buffers, device buffers, ragged sizes, and over many consecutive scans."""
    import ctypes as C
    import torch
    from mins_b200 import capi
    # "auto": ml_scan_update picks the path by scan size and search depth (here: front + kNN kernel, then the rows kernel);
    # "one_kernel": k_scan_tile forced (MINS_SCAN_TILE=2 is read when the handle is created), on ragged sizes down to one point
    if path == "one_kernel":
        monkeypatch.setenv("MINS_SCAN_TILE", "2")
    w = workload("G-vlp")
    op = w["op"]
    m = make_map(max_map_points=1 << 19)
    m.Build(w["map"])
    hx = w["total_hx"]
    T = np.ascontiguousarray(w["T_LtoM"], np.float32)
    P = np.ascontiguousarray(w["P"], np.float64)

    def one_call(ptr, stride, n):
        HtH, Htr, cnt = np.zeros((hx, hx)), np.zeros(hx), capi.ml_counters()
        rc = m.lib.ml_scan_update(m.h, ptr, stride, n, w["t_rel"], op.min_range, op.max_range, 0.0, T.ctypes.data,
                                  C.byref(w["consts"]), P.ctypes.data, HtH.ctypes.data, Htr.ctypes.data, C.byref(cnt))
        assert rc in (0, 1)
        return HtH, Htr, cnt.as_dict(), m.scan(0), m.scan(1), m.scan(2)

    raw = np.ascontiguousarray(w["raw"], np.float32)
    # make a third of the points fall outside the range gate so that the compaction is non-trivial in every tile
    raw = raw.copy()
    raw[::3, :3] *= 1000.0
    for n in (len(raw), 2048 * 3 + 17, 2047, 1):
        sub = np.ascontiguousarray(raw[:n])
        # reference state: the multi-call path
        kept = m.feed_scan(sub, w["t_rel"], op)
        m.transform_to_map(T)
        idx0, d20, ok0 = m.get_neighbors(op.map_ngbr_num, op.map_ngbr_max_d)
        rc0, HtH0, Htr0, cnt0 = m.linsys(w["consts"], P)
        s0 = (m.scan(0), m.scan(1), m.scan(2))
        assert len(s0[0]) == kept == len(po.range_filter(sub, w["t_rel"], op.max_range, op.min_range))
        pinned = torch.from_numpy(sub).pin_memory()
        dev = torch.from_numpy(sub).cuda()
        xyz3 = np.ascontiguousarray(sub[:, :3])
        first = None
        for ptr, stride in ((sub.ctypes.data, 4), (pinned.data_ptr(), 4), (dev.data_ptr(), 4), (xyz3.ctypes.data, 3)):
            for _ in range(3):
                HtH, Htr, cnt, a, b, c = one_call(ptr, stride, n)
                assert np.array_equal(a, s0[0]) and np.array_equal(b, s0[1]) and np.array_equal(c, s0[2])
                # the one-kernel path adds the same per-point rows in another (fixed) order than the multi-call path's rows
                # kernel: integer counters identical, floating-point sums equal to rounding ...
                ref = cnt0.as_dict()
                for key, v in cnt.items():
                    assert v == ref[key] if isinstance(ref[key], int) else abs(v - ref[key]) <= 1e-12 * max(1.0, abs(ref[key])), key
                assert rel(HtH, HtH0) < 1e-12 and rel(Htr, Htr0) < 1e-12
                # ... and bit-identical from call to call, whatever kind of buffer the scan arrives in
                if first is None:
                    first = (HtH, Htr, cnt)
                assert cnt == first[2] and np.array_equal(HtH, first[0]) and np.array_equal(Htr, first[1])
            idx, d2, ok = m.get_neighbors(op.map_ngbr_num, op.map_ngbr_max_d)
            assert np.array_equal(idx, idx0) and np.array_equal(d2, d20) and np.array_equal(ok, ok0)
