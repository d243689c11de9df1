"""GPU parity at the BASELINE.json configurations, through the C ABI.

 * G-64 (config[2], 131 072 rays vs 1 000 000 map points): EVERY query's k nearest squared distances against the
   reference's own ikd-Tree compiled unmodified (oracle/_ref/libikd_ref.so travels to the GPU box), neighbour coordinates
   wherever the answer is unique, and the reduced system against the oracle rows driven by the REFERENCE TREE's neighbours.
 * the same with raw_do_downsample (leaf 0.3, the reference's default) at G-64.
 * the KAIST option set (mins/config/kaist/config_lidar.yaml:11-29) on the 2 x VLP-16 shape (config[1]).
 * re-anchor with down-sampling held to the voxel contract; rank-deficient neighbour sets; chi2 gate for k != 5.
"""
import numpy as np
import pytest

from conftest import check_voxel_contract, rel

pytestmark = pytest.mark.gpu
TOL = 1e-9


@pytest.fixture()
def make_map(cuda_ok):
    from mins_b200.lidar import LidarMap
    made = []

    def mk(**kw):
        m = LidarMap(**kw)
        made.append(m)
        return m

    yield mk
    for m in made:
        m.close()


def dist2_f32(a, b):
    """KD_TREE::calc_dist (ikd_Tree.cpp:1406-1410) in fp32 without contraction"""
    d = a[..., :3].astype(np.float32) - b[..., :3].astype(np.float32)
    return ((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]).astype(np.float32) + d[..., 2] * d[..., 2]).astype(np.float32)


def against_reference_tree(po, m, w, op, HtH, Htr, cnt, threads=16):
    """Every query of the resident scan against the reference's KD_TREE::Nearest_Search; then the oracle rows driven by the
    reference tree's neighbours against the device's reduced system. Returns (#queries, #queries whose answer is not unique)."""
    k, max_d = op.map_ngbr_num, op.map_ngbr_max_d
    sc, sm = m.scan(0), m.scan(1)
    idx, d2, ok = m.get_neighbors(k, max_d)
    tree = po.RefIkd(op.map_downsample_size)
    tree.build(w["map"])
    nb, rd2, found = tree.knn(sm, k, threads=threads)
    assert (found == k).all()
    rok = ~(rd2[:, k - 1].astype(np.float64) > max_d)            # LidarHelper.cpp:280-285 (squared vs unsquared: kept)
    assert np.array_equal(ok, rok)
    assert np.array_equal(d2[ok], rd2[ok])                        # all k squared distances of every accepted query, bit for bit
    # queries get_neighbors rejects: the device reports candidates within the radius only (-1 / inf beyond)
    inside = rd2 <= np.float32(max_d)
    assert np.array_equal(np.where(inside, rd2, np.float32(np.inf))[~ok], d2[~ok])
    assert ((idx >= 0) == inside).all()
    # neighbour identity: the device's map points are the reference's, coordinate for coordinate, wherever the answer is
    # unique; where two map points are exactly equidistant the tree's pick depends on its traversal (not on the point set)
    mine = w["map"][np.where(idx >= 0, idx, 0)]
    same = (mine[..., :3] == nb[..., :3]).all(-1) | ~inside
    odd = ~same.all(1)
    # a differing pick must be an exact tie: same distance to the query from both points
    assert np.array_equal(dist2_f32(mine[odd], sm[odd][:, None, :])[inside[odd]], rd2[odd][inside[odd]])
    assert odd.sum() <= max(2, len(sm) // 5000), int(odd.sum())
    nb_used = nb.copy()
    nb_used[odd] = mine[odd]                                      # both are exact answers; use the device's pick for those few
    ref = po.scan_linsys_nb(w["consts"], w["P"], sc, nb_used, rok, threads=threads)
    gc, oc = cnt.as_dict(), ref["cnt"].as_dict()
    for key in ("n_points", "n_knn_fail", "n_plane_fail", "n_badmatch", "n_chi_rej", "n_acc"):
        assert gc[key] == oc[key], key
    assert gc["sum_chi"] == pytest.approx(oc["sum_chi"], rel=TOL) and gc["sum_res2"] == pytest.approx(oc["sum_res2"], rel=TOL)
    assert rel(HtH, ref["HtH"]) <= TOL and rel(Htr, ref["Htr"]) <= TOL
    return len(sm), int(odd.sum())


def test_g64_every_query_against_the_reference_tree(make_map, po, workload):
    w = workload("G-64")
    op = w["op"]
    n_rays, M = w["n_rays"], len(w["map"])
    assert n_rays == 131072 and M == 1_000_000 and not op.raw_do_downsample
    m = make_map(max_scan_points=n_rays, max_map_points=M + 2 * n_rays)
    m.Build(w["map"])
    rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert rc == 0
    sc = po.range_filter(w["raw"], w["t_rel"], op.max_range, op.min_range)
    assert np.array_equal(m.scan(0), sc) and np.array_equal(m.scan(1), po.transform_f32(sc, w["T_LtoM"]))
    n, odd = against_reference_tree(po, m, w, op, HtH, Htr, cnt)
    assert n == len(sc) > 120000 and cnt.n_acc > 0.4 * n


def test_g64_one_kernel_path_against_the_two_kernel_path(make_map, workload, monkeypatch):
    """k_scan_tile (one thread per ray, tile.cuh) against k_front_knn + k_rows_reduce (one warp per cell) on the headline scan: same
    clouds, same neighbours (the rows export re-reads them), same counters; the reduced system equal to rounding (the two add the same
    rows in different fixed orders) and bit-identical from call to call on either path. The KAIST option set on the same scan
    (7 levels deep) must take the two-kernel path by itself."""
    w = workload("G-64")
    op = w["op"]
    n_rays, M = w["n_rays"], len(w["map"])
    res = {}
    for path, env in (("one", "2"), ("one_lazy", "2"), ("two", "0")):
        monkeypatch.setenv("MINS_SCAN_TILE", env)
        # "one": k_scan_tile also stores the neighbours it found (MINS_TILE_NBR=1), so the rows export below reads the one-thread-per-
        # query walk's own neighbours; "one_lazy": the default, the export regenerates them with the multi-call search
        monkeypatch.setenv("MINS_TILE_NBR", "1" if path == "one" else "0")
        m = make_map(max_scan_points=n_rays, max_map_points=M + 2 * n_rays)
        m.Build(w["map"])
        l0 = m.lib.ml_launch_count(m.h)
        rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
        launches = m.lib.ml_launch_count(m.h) - l0
        rc2, HtH2, Htr2, cnt2 = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
        assert rc == 0 and rc2 == 0 and launches == (2 if path == "two" else 1)
        assert np.array_equal(HtH, HtH2) and np.array_equal(Htr, Htr2) and cnt.as_dict() == cnt2.as_dict()
        rows, _ = m.point_rows(want_h=False)
        res[path] = (HtH, Htr, cnt.as_dict(), m.scan(0), m.scan(1), m.scan(2), rows)
    monkeypatch.delenv("MINS_SCAN_TILE")
    monkeypatch.delenv("MINS_TILE_NBR")
    a, b, lazy = res["one"], res["two"], res["one_lazy"]
    assert np.array_equal(a[0], lazy[0]) and np.array_equal(a[1], lazy[1]) and a[2] == lazy[2]   # storing the neighbours changes nothing
    for k in (3, 4, 5):
        assert np.array_equal(a[k], b[k]) and np.array_equal(lazy[k], b[k])
    for r in (a[6], lazy[6]):
        assert np.array_equal(r["status"], b[6]["status"]) and np.array_equal(r["plane"], b[6]["plane"])
        assert np.array_equal(r["chi2"], b[6]["chi2"], equal_nan=True)
    for key, v in a[2].items():
        assert v == b[2][key] if isinstance(v, int) else abs(v - b[2][key]) <= 1e-12 * max(1.0, abs(v)), key
    assert rel(a[0], b[0]) < 1e-12 and rel(a[1], b[1]) < 1e-12
    # automatic choice: the headline scan takes one launch, the same scan with the KAIST search radius two
    m = make_map(max_scan_points=n_rays, max_map_points=M + 2 * n_rays)
    m.Build(w["map"])
    for max_d, want in ((op.map_ngbr_max_d, 1), (10.0, 2)):
        c = type(w["consts"]).from_buffer_copy(w["consts"])
        c.map_ngbr_max_d = max_d
        l0 = m.lib.ml_launch_count(m.h)
        rc, _, _, _ = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], c, w["P"])
        assert rc == 0 and m.lib.ml_launch_count(m.h) - l0 == want


def test_one_kernel_path_with_more_tiles_than_ctas(make_map, workload, monkeypatch):
    """A scan of 262 144 rays is 586 tiles for at most 296 resident CTAs: every CTA of k_scan_tile walks two tiles (second TMA copy
    into the reused staging area, look-back across rounds, the CTA's running system over both). Checked against the two-kernel path
    on the same scan: clouds and counters identical, reduced system equal to rounding, repeatable bit for bit."""
    w = workload("G-64")
    op = w["op"]
    raw = np.ascontiguousarray(np.concatenate([w["raw"], w["raw"][::-1]]))
    n_rays, M = len(raw), len(w["map"])
    assert n_rays == 262144
    res = {}
    for path, env in (("one", "2"), ("two", "0")):
        monkeypatch.setenv("MINS_SCAN_TILE", env)
        m = make_map(max_scan_points=n_rays, max_map_points=M + 8)
        m.Build(w["map"])
        rc, HtH, Htr, cnt = m.scan_update(raw, w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
        rc2, HtH2, Htr2, cnt2 = m.scan_update(raw, w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
        assert rc == 0 and rc2 == 0
        assert np.array_equal(HtH, HtH2) and np.array_equal(Htr, Htr2) and cnt.as_dict() == cnt2.as_dict()
        res[path] = (HtH, Htr, cnt.as_dict(), m.scan(0), m.scan(1), m.scan(2))
    monkeypatch.delenv("MINS_SCAN_TILE")
    a, b = res["one"], res["two"]
    assert len(a[3]) > 240000
    for k in (3, 4, 5):
        assert np.array_equal(a[k], b[k])
    for key, v in a[2].items():
        assert v == b[2][key] if isinstance(v, int) else abs(v - b[2][key]) <= 1e-12 * max(1.0, abs(v)), key
    assert rel(a[0], b[0]) < 1e-12 and rel(a[1], b[1]) < 1e-12


def test_g64_mode_b_downsampled_scan(make_map, po, workload):
    """raw_do_downsample: true, leaf 0.3 (mins/config/simulation/config_lidar.yaml:14-15) at the headline size"""
    w = workload("G-64", downsample=True)
    op = w["op"]
    assert op.raw_do_downsample and op.raw_downsample_size == 0.3
    n_rays, M = w["n_rays"], len(w["map"])
    m = make_map(max_scan_points=n_rays, max_map_points=M + 2 * n_rays)
    m.Build(w["map"])
    rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert rc == 0
    in_range = po.range_filter(w["raw"], w["t_rel"], op.max_range, op.min_range)
    assert np.array_equal(m.scan(2), in_range)                    # pointcloud_original keeps every in-range point
    out = m.scan(0)
    counts = check_voxel_contract(in_range, out, op.raw_downsample_size, po)
    assert 0.05 * len(in_range) < len(out) < 0.5 * len(in_range) and counts.max() > 20
    assert np.array_equal(m.scan(1), po.transform_f32(out, w["T_LtoM"]))
    n, odd = against_reference_tree(po, m, w, op, HtH, Htr, cnt)
    assert n == len(out) and cnt.n_acc > 0.2 * n  # the far half of the scan fails the condition gate (LidarHelper.cpp:310)


@pytest.mark.parametrize("downsample", [True, False])
def test_kaist_option_set(make_map, po, workload, downsample):
    """leaf 2.0, map voxel 0.5, map_ngbr_max_d 10 (radius sqrt(10) m: the cube walk reaches 7 cells), raw/map noise 0.1,
    max_range 30 - on the 2 x VLP-16 scan. Oracle: brute-force kNN + restated rows on the device's own query cloud."""
    w = workload("G-vlp", downsample=downsample, options="kaist")
    op = w["op"]
    assert op.map_ngbr_max_d == 10.0 and op.raw_downsample_size == 2.0 and op.map_downsample_size == 0.5
    # thin the map out ahead of the sensor (1 point in 40 beyond x = 3 m) so that part of the scan finds its 5th neighbour
    # metres away: cube walks to the last level, and queries beyond the gate (k-th squared distance > 10, LidarHelper.cpp:284)
    mp = w["map"]
    mp = np.ascontiguousarray(mp[(mp[:, 0] < 3.0) | (np.arange(len(mp)) % 40 == 0)])
    m = make_map(max_scan_points=1 << 16, max_map_points=1 << 18)
    m.Build(mp)
    rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], op, w["T_LtoM"], w["consts"], w["P"])
    assert rc == 0
    in_range = po.range_filter(w["raw"], w["t_rel"], op.max_range, op.min_range)
    assert np.array_equal(m.scan(2), in_range)
    sc, sm = m.scan(0), m.scan(1)
    if downsample:
        check_voxel_contract(in_range, sc, op.raw_downsample_size, po)
    else:
        assert np.array_equal(sc, in_range)
    assert np.array_equal(sm, po.transform_f32(sc, w["T_LtoM"]))
    rows, h = m.point_rows()                                       # the export must follow the update directly
    idx, d2, ok = m.get_neighbors(op.map_ngbr_num, op.map_ngbr_max_d)
    oidx, od2, ook = po.knn_brute(mp, sm, op.map_ngbr_num, np.float32(op.map_ngbr_max_d))
    assert np.array_equal(ok, ook) and np.array_equal(idx, oidx) and np.array_equal(d2, od2)
    assert (od2[ook][:, -1] > 4.0).sum() > 20 and (~ook).sum() > 20   # 5th neighbour beyond 2 m; beyond the gate
    ref = po.scan_linsys(w["consts"], w["P"], mp, sc, sm)
    assert np.array_equal(rows["status"], ref["rows"]["status"])
    assert cnt.n_acc == ref["cnt"].n_acc > 0.1 * cnt.n_points
    both = rows["status"] == 0
    assert rel(h[both], ref["h"][both]) <= TOL
    assert rel(HtH, ref["HtH"]) <= TOL and rel(Htr, ref["Htr"]) <= TOL
    # the map side of the option set: register the scan at 0.5 m voxels, re-anchor with decay 30 s / 30 m
    n_live = m.register_scan(op)
    ref_map = po.map_insert(mp, po.transform_f32(in_range, w["T_LtoM"]), True, op.map_downsample_size)
    assert n_live == len(ref_map) and np.array_equal(m.flatten(), ref_map)


def test_reanchor_with_downsampling_meets_the_voxel_contract(make_map, po, workload):
    """LidarHelper::propagate_map_to_newest_clone with map_do_downsample (LidarHelper.cpp:209-266): identical voxel
    membership and order, centroids within 4 ulp sqrt(n), then the decay filters and the fp64 transform of the survivors."""
    from mins_b200.lidar import OptionsLidar
    from mins_b200.synth import exp_so3
    w = workload("G-vlp")
    op = OptionsLidar(map_do_downsample=True, map_downsample_size=0.6, map_decay_time=3.0, map_decay_dist=40.0)
    T = np.eye(4)
    T[:3, :3] = exp_so3(np.array([0.01, 0.02, -0.05]))
    T[:3, 3] = [-1.0, 0.05, 0.02]
    now = 5.5
    m = make_map(max_scan_points=1 << 16, max_map_points=1 << 18)
    m.Build(w["map"])
    n = m.propagate_map_to_newest_clone(T, now, op)
    out = m.flatten()
    ref_full = po.reanchor(w["map"], T, now, op.map_decay_time, op.map_decay_dist, True, op.map_downsample_size)

    def post(c):  # LidarHelper.cpp:219-226 on the oracle's centroids, then :261
        keep = ~(c[:, 3].astype(np.float64) + op.map_decay_time < now)
        keep &= ~(np.sqrt((c[:, 0] * c[:, 0] + (c[:, 1] * c[:, 1] + c[:, 2] * c[:, 2]).astype(np.float32)).astype(np.float32)) > op.map_decay_dist)
        q = c[keep].copy()
        q[:, :3] = (c[keep, :3].astype(np.float64) @ T[:3, :3].T + T[:3, 3]).astype(np.float32)
        return q, keep

    counts = check_voxel_contract(w["map"], out, op.map_downsample_size, po, post=post, slack=2.0)
    assert n == len(out) == len(ref_full) and 0 < n < len(w["map"]) and counts.max() >= 4
    assert np.abs(out - ref_full).max() <= 1e-4                    # and the oracle's own end-to-end result agrees


def test_rank_deficient_neighbour_sets(make_map, po):
    """compute_plane on degenerate neighbourhoods (LidarHelper.cpp:305-312): the reference's sigma_max / sigma_min is huge or
    inf (> plane_max_condi: rejected) for collinear / coplanar-through-the-origin / coincident neighbours; the device takes
    the same decision from the eigenvalues of A^T A (csrc/rows.cuh, `!(lmin > 0)`)."""
    from mins_b200 import synth
    from mins_b200.lidar import OptionsLidar
    w = synth.make_workload("G-small")
    op = OptionsLidar(raw_do_downsample=False, map_ngbr_max_d=4.0)
    import copy
    c = copy.copy(w["consts"])
    c.map_ngbr_max_d = 4.0
    t = np.arange(5, dtype=np.float32)[:, None]
    cases = {
        "collinear through the origin": t * np.float32([0.3, 0.3, 0.0]) + np.float32([0.3, 0.3, 0.0]),
        "collinear off the origin": t * np.float32([0.25, 0.0, 0.0]) + np.float32([3.0, 2.0, 0.5]),
        "coplanar with the origin": np.float32([[2, 0, 0], [2.3, 0, 0], [2, 0.3, 0], [2.3, 0.3, 0], [2.15, 0.15, 0]]),
        "axis-aligned wall (well conditioned reference)": np.float32([[3, 0, 0], [3, 0.3, 0], [3, 0, 0.3], [3, 0.3, 0.3], [3, 0.15, 0.6]]),
    }
    for name, nbp in cases.items():
        mp = np.zeros((5, 4), np.float32)
        mp[:, :3] = nbp
        q = np.zeros((3, 4), np.float32)
        q[:, :3] = nbp.mean(0) + np.float32([[0.01, 0.02, 0.03], [0.0, 0.0, 0.05], [-0.02, 0.01, -0.04]])
        m = make_map(max_scan_points=1024, max_map_points=1024)
        m.Build(mp)
        m.set_scan(q)
        m.transform_to_map(np.eye(4, dtype=np.float32))
        rc, HtH, Htr, cnt = m.linsys(c, w["P"])
        ref = po.scan_linsys(c, w["P"], mp, q, q)
        rows, _ = m.point_rows(want_h=False)
        assert ref["cnt"].n_knn_fail == 0, name
        assert np.array_equal(rows["status"], ref["rows"]["status"]), (name, rows["status"], ref["rows"]["status"])
        ok_plane, plane, cond = po.compute_plane(mp)
        if "wall" in name:
            assert ok_plane and (rows["status"] != 2).all()
        else:
            assert not ok_plane and not (cond <= op.plane_max_condi) and (rows["status"] == 2).all(), (name, cond)
            assert cnt.n_plane_fail == 3 and cnt.n_acc == 0
        m.close()


@pytest.mark.parametrize("k", [3, 4, 7, 10])
def test_chi2_gate_uses_the_projected_row_count(make_map, po, workload, k):
    """UpdaterStatistics::get_chi2 compares with chi_squared_table[res.rows()], res.rows() = k + 1 - 3 after the nullspace
    projection (UpdaterStatistics.cpp:110-111, UpdaterLidar.cpp:301,362): 1, 2, 5 and 8 DoF here, with chi2_mult chosen so
    that the gate splits the points. Decisions against the oracle AND against scipy's quantile on the device's own chi2."""
    import copy
    from scipy.stats import chi2 as chi2_dist
    w = workload("G-small", k=k)
    m = make_map(max_scan_points=1 << 14, max_map_points=1 << 16)
    m.Build(w["map"])
    rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], w["op"], w["T_LtoM"], w["consts"], w["P"])
    rows, _ = m.point_rows(want_h=False)
    tested = (rows["status"] == 0) | (rows["status"] == 4)
    q = float(chi2_dist.ppf(0.95, k - 2))
    c2 = copy.copy(w["consts"])
    v = np.sort(rows["chi2"][tested])
    c2.chi2_mult = float(0.5 * (v[len(v) // 2] + v[len(v) // 2 + 1])) / q   # half of the tested points must now fail, nobody on the gate
    rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], w["op"], w["T_LtoM"], c2, w["P"])
    rows, _ = m.point_rows(want_h=False)
    ref = po.scan_linsys(c2, w["P"], w["map"], m.scan(0), m.scan(1))
    assert np.array_equal(rows["status"], ref["rows"]["status"])
    assert cnt.n_chi_rej > 0.3 * tested.sum() and cnt.n_acc > 0.3 * tested.sum()
    assert np.array_equal(rows["status"][tested] == 0, rows["chi2"][tested] < c2.chi2_mult * q)
    assert rel(HtH, ref["HtH"]) <= TOL and rel(Htr, ref["Htr"]) <= TOL


@pytest.mark.parametrize("variant,k", [("plain", 5), ("calib", 5), ("overlap", 7), ("plain", 3)])
def test_residual_statistics_first_moment(make_map, po, workload, variant, k):
    """ml_counters.sum_res: res_stat->add_stat(res) of UpdaterStatistics::Chi2Check (UpdaterStatistics.cpp:54-57) sums the
    projected residuals r' themselves - a number that depends on the basis StateHelper::Givens_Rotation (StateHelper.cpp:660-680)
    leaves, so the device runs that very rotation sequence (statistics on); off, the field stays 0 and nothing else changes."""
    w = workload("G-small", variant=variant, k=k)
    m = make_map(max_scan_points=1 << 14, max_map_points=1 << 16)
    m.Build(w["map"])
    rc, HtH0, Htr0, cnt0 = m.scan_update(w["raw"], w["t_rel"], w["op"], w["T_LtoM"], w["consts"], w["P"])
    assert rc == 0 and cnt0.sum_res == 0.0
    m.set_stats(True)
    rc, HtH, Htr, cnt = m.scan_update(w["raw"], w["t_rel"], w["op"], w["T_LtoM"], w["consts"], w["P"])
    ref = po.scan_linsys(w["consts"], w["P"], w["map"], m.scan(0), m.scan(1))
    oc = ref["cnt"]
    assert cnt.n_acc == oc.n_acc == cnt0.n_acc
    scale = np.sqrt(oc.sum_res2 * (cnt.n_acc + cnt.n_chi_rej) * (k - 2))       # Cauchy-Schwarz bound of |sum r'|
    assert abs(cnt.sum_res - oc.sum_res) <= 1e-9 * scale and abs(oc.sum_res) > 1e-6 * scale
    assert cnt.sum_res2 == pytest.approx(oc.sum_res2, rel=TOL) and cnt.sum_chi == pytest.approx(oc.sum_chi, rel=TOL)
    assert rel(HtH, ref["HtH"]) <= TOL and rel(HtH, HtH0) <= 1e-12 and rel(Htr, Htr0) <= 1e-11
    rows, h = m.point_rows()                                                       # the rows are already there
    assert np.array_equal(rows["status"], ref["rows"]["status"])
