"""Cross-checks of the oracle's hand-written small dense algebra against numpy.linalg (CPU only).

The reference uses Eigen 3.3.7 (JacobiSVD, ColPivHouseholderQR, makeGivens) which is not in this image; the oracle restates
them (oracle/lidar_oracle.cpp). These tests show the restatement computes what those decompositions define, to <= 1e-12
relative, by an independent route (numpy SVD / lstsq / QR), and that the per-point closed forms the CUDA kernels use are
algebraically the reference's Givens + 3x3-inverse route (UpdaterLidar.cpp:352-368, StateHelper.cpp:645-680,
UpdaterStatistics.cpp:94-117).
"""
import ctypes as C

import numpy as np
import pytest

from conftest import rel


def plane_neighbours(rng, k, noise=0.02, offset=(8.0, -3.0, 1.5)):
    n = rng.normal(size=3)
    n /= np.linalg.norm(n)
    u = np.cross(n, [1, 0, 0.3])
    u /= np.linalg.norm(u)
    v = np.cross(n, u)
    uv = rng.uniform(-0.5, 0.5, (k, 2))
    p = np.asarray(offset) + uv[:, :1] * u + uv[:, 1:] * v + rng.normal(0, noise, (k, 1)) * n
    nb = np.zeros((k, 4), np.float32)
    nb[:, :3] = p
    return nb


@pytest.mark.parametrize("k", [3, 5, 7, 10])
def test_compute_plane_matches_numpy_svd_and_lstsq(po, k):
    """LidarHelper.cpp:291-328: cond(A) by SVD, abc = argmin ||A abc + d 1||, d = mean range, gates."""
    rng = np.random.default_rng(k)
    n_ok = 0
    for trial in range(300):
        nb = plane_neighbours(rng, k, noise=rng.choice([0.0, 0.01, 0.05, 0.3]))
        ok, plane, cond = po.compute_plane(nb, 200.0, 0.1)
        A = nb[:, :3].astype(np.float64)
        sv = np.linalg.svd(A, compute_uv=False)
        assert cond == pytest.approx(sv[0] / sv[-1], rel=1e-10)
        d = np.linalg.norm(A, axis=1).mean()
        if sv[0] / sv[-1] > 200.0:
            assert not ok
            continue
        abc = np.linalg.lstsq(A, -d * np.ones(k), rcond=None)[0]
        assert np.abs(plane[:3] - abc).max() <= 1e-10 * max(np.abs(abc).max(), 1e-12) * (sv[0] / sv[-1])
        assert plane[3] == pytest.approx(d, rel=1e-14)
        p2p = np.abs(A @ abc + d) / np.linalg.norm(abc)
        if abs(p2p.max() - 0.1) > 1e-9:
            assert ok == bool(p2p.max() <= 0.1)
        n_ok += ok
    assert n_ok > 50


def numpy_point(c, P, pfinL, nb):
    """independent numpy route for one point: explicit H_f, H, res; nullspace by full QR; chi2 by solve."""
    k, hx = c.map_ngbr_num, c.total_hx
    g = lambda name, shape: np.array(getattr(c, name)).reshape(shape)
    RLtoM_est, pLinM_est = g("RLtoM_est", (3, 3)), g("pLinM_est", 3)
    RLtoM_fej, pLinM_fej = g("RLtoM_fej", (3, 3)), g("pLinM_fej", 3)
    RGtoM, RItoG, RLtoI, RItoL = g("RGtoM_fej", (3, 3)), g("RItoG_fej", (3, 3)), g("RLtoI_fej", (3, 3)), g("RItoL_fej", (3, 3))
    RGtoA, RLtoG = g("RGtoA_fej", (3, 3)), g("RLtoG_fej", (3, 3))
    pIinL, pLinI, pIinG, pAinG = g("pIinL_fej", 3), g("pLinI_fej", 3), g("pIinG_fej", 3), g("pAinG_fej", 3)
    sk = lambda w: np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    A = nb[:, :3].astype(np.float64)
    d = np.linalg.norm(A, axis=1).mean()
    abc = np.linalg.lstsq(A, -d * np.ones(k), rcond=None)[0]
    pL = pfinL.astype(np.float64)
    pM_est, pM_fej = pLinM_est + RLtoM_est @ pL, pLinM_fej + RLtoM_fej @ pL
    res = np.concatenate([-(A @ abc + d), [-(abc @ pM_est + d)]])
    Hf = np.vstack([A, pM_fej])
    dI = np.hstack([-RGtoM @ RItoG @ sk(RLtoI @ (pL - pIinL)), RGtoM])
    dA = np.hstack([RItoL @ sk(RGtoA @ (pIinG + RItoG @ pLinI - pAinG + RLtoG @ pL)), -RGtoM])
    dC = np.hstack([sk(RGtoM @ (pIinG - pAinG)) - sk(RLtoM_fej @ (pIinL - pL)) + RLtoM_fej @ sk(pIinL - pL), np.eye(3) - RLtoM_fej])
    h = np.zeros(hx)
    for j in range(c.nI):
        B = c.dTdxI[j]
        h[B.col:B.col + B.width] += (abc @ dI) @ np.array(B.D).reshape(6, 6)[:, :B.width]
    for j in range(c.nA):
        B = c.dTdxA[j]
        h[B.col:B.col + B.width] += (abc @ dA) @ np.array(B.D).reshape(6, 6)[:, :B.width]
    if c.calib_col >= 0:
        h[c.calib_col:c.calib_col + 6] += abc @ dC
    intr = 0.0
    if not (c.at_clone_m and c.at_clone_a):
        HI = abc @ dI
        intr = np.sqrt(c.ie_o * HI[:3] @ HI[:3] + c.ie_p * HI[3:] @ HI[3:])
    nn = np.linalg.norm(abc)
    raw_n, map_n = nn * c.raw_noise + intr, nn * c.map_noise + intr
    H = np.zeros((k + 1, hx))
    H[k] = h / raw_n
    Hf[k] /= raw_n
    Hf[:k] /= map_n
    res[k] /= raw_n
    res[:k] /= map_n
    Q, _ = np.linalg.qr(Hf, mode="complete")
    N = Q[:, 3:]                                    # left nullspace of H_f
    Hp, rp = N.T @ H, N.T @ res
    S = Hp @ P @ Hp.T + np.eye(k - 2)
    chi2 = rp @ np.linalg.solve(S, rp)
    return dict(plane=np.append(abc, d), Hp=Hp, rp=rp, chi2=chi2, h=H[k], res=res[k], raw_n=raw_n)


@pytest.mark.parametrize("variant,k", [("plain", 5), ("calib_dt", 5), ("overlap", 7), ("at_clone", 3), ("calib", 10)])
def test_point_rows_match_independent_numpy_route(po, workload, variant, k):
    from mins_b200 import capi
    w = workload("G-small", variant=variant, k=k)
    c, P, op = w["consts"], w["P"], w["op"]
    sc = po.range_filter(w["raw"], w["t_rel"], op.max_range, op.min_range)
    sm = po.transform_f32(sc, w["T_LtoM"])
    idx, d2, ok = po.knn_brute(w["map"], sm, k, np.float32(op.map_ngbr_max_d))
    L = po.lib()
    hx = c.total_hx
    checked = 0
    for i in np.where(ok)[0][::7][:150]:
        nb = np.ascontiguousarray(w["map"][idx[i]])
        Hp, rp = np.zeros((k - 2, hx)), np.zeros(k - 2)
        row, hrow = capi.ml_point_row(), np.zeros(hx)
        st = L.orc_point(C.byref(c), P.ctypes.data, sc[i].ctypes.data, nb.ctypes.data, Hp.ctypes.data, rp.ctypes.data,
                         C.byref(row), hrow.ctypes.data)
        if st == 2:
            continue
        ref = numpy_point(c, P, sc[i, :3], nb)
        assert rel(np.array(row.plane), ref["plane"]) <= 1e-9
        if st == 3:
            continue
        checked += 1
        # the projection basis is not unique; its invariants are: H'^T H', H'^T r', ||r'||, chi2
        assert rel(Hp.T @ Hp, ref["Hp"].T @ ref["Hp"]) <= 1e-9
        assert rel(Hp.T @ rp, ref["Hp"].T @ ref["rp"]) <= 1e-8
        assert rp @ rp == pytest.approx(ref["rp"] @ ref["rp"], rel=1e-9)
        assert row.chi2 == pytest.approx(ref["chi2"], rel=1e-9)
        assert rel(hrow, ref["h"]) <= 1e-9 and row.res == pytest.approx(ref["res"], rel=1e-9, abs=1e-12)
        assert row.raw_n == pytest.approx(ref["raw_n"], rel=1e-12)
        # rank-one structure used by the CUDA path (DESIGN.md "closed forms"): H' = v h^T
        assert np.linalg.matrix_rank(Hp, tol=1e-9 * np.abs(Hp).max()) == 1
        assert rel(Hp.T @ Hp, row.w * np.outer(hrow, hrow)) <= 1e-9
        assert rel(Hp.T @ rp, row.vr * hrow) <= 1e-8
        s = hrow @ P @ hrow
        assert row.chi2 == pytest.approx(rp @ rp - s * row.vr ** 2 / (1 + s * row.w), rel=1e-9)
    assert checked > 30


def test_scan_linsys_stack_and_compression_are_consistent(po, workload):
    """UpdaterLidar.cpp:376-397 + StateHelper.cpp:631-643: Hc^T Hc = H_big^T H_big, upper-trapezoidal Hc, and
    HtH = sum_i w_i h_i h_i^T over accepted points."""
    w = workload("G-small")
    op = w["op"]
    sc = po.range_filter(w["raw"], w["t_rel"], op.max_range, op.min_range)
    sm = po.transform_f32(sc, w["T_LtoM"])
    ref = po.scan_linsys(w["consts"], w["P"], w["map"], sc, sm, want_compress=True)
    acc = ref["rows"]["status"] == 0
    assert ref["total_res"] == 3 * acc.sum() and ref["cnt"].n_acc == acc.sum()
    HtH = (ref["h"][acc] * ref["rows"]["w"][acc, None]).T @ ref["h"][acc]
    Htr = ref["h"][acc].T @ ref["rows"]["vr"][acc]
    assert rel(HtH, ref["HtH"]) <= 1e-10 and rel(Htr, ref["Htr"]) <= 1e-9
    Hc, rc = ref["Hc"], ref["rc"]
    assert Hc.shape == (48, 48) and np.abs(np.tril(Hc, -1)).max() <= 1e-12 * np.abs(Hc).max()
    assert rel(Hc.T @ Hc, ref["HtH"]) <= 1e-10 and rel(Hc.T @ rc, ref["Htr"]) <= 1e-9


def test_ekf_update_matches_numpy(po):
    rng = np.random.default_rng(11)
    n, m = 30, 12
    L = rng.normal(size=(n, n))
    P = L @ L.T / n + np.eye(n)
    H, r = rng.normal(size=(m, n)), rng.normal(size=m)
    dx, Pn = po.ekf_update(P, H, r)
    S = H @ P @ H.T + np.eye(m)
    K = P @ H.T @ np.linalg.inv(S)
    assert rel(dx, K @ r) <= 1e-11 and rel(Pn, P - K @ H @ P) <= 1e-11


def test_voxel_downsample_matches_numpy_restatement(po):
    """PCL VoxelGrid semantics (documented algorithm): centroid of all four fields per voxel, ascending voxel index."""
    rng = np.random.default_rng(12)
    pts = np.zeros((4000, 4), np.float32)
    pts[:, :3] = rng.uniform(-7, 9, (4000, 3))
    pts[:, 3] = rng.uniform(0, 3, 4000)
    leaf = np.float32(0.8)
    out, vid = po.voxel_downsample(pts, float(leaf))
    inv = np.float32(1.0) / leaf
    ijk = np.floor(pts[:, :3] * inv).astype(np.int64)
    mn = np.floor(pts[:, :3].min(0) * inv).astype(np.int64)
    dv = np.floor(pts[:, :3].max(0) * inv).astype(np.int64) - mn + 1
    lin = (ijk[:, 0] - mn[0]) + (ijk[:, 1] - mn[1]) * dv[0] + (ijk[:, 2] - mn[2]) * dv[0] * dv[1]
    u = np.unique(lin)
    assert np.array_equal(u, vid.astype(np.int64))
    cen = np.stack([pts[lin == v].astype(np.float64).mean(0) for v in u])
    assert np.abs(out - cen).max() <= 2e-5
    # leaf so small that the voxel index overflows int32: PCL returns the input unchanged
    big = pts.copy()
    big[:, :3] *= 100
    out2, _ = po.voxel_downsample(big, 0.01)  # 1.6e5 voxels per axis: 4e15 > INT32_MAX
    assert np.array_equal(out2, big)


def test_reanchor_matches_numpy(po):
    rng = np.random.default_rng(13)
    m = np.zeros((3000, 4), np.float32)
    m[:, :3] = rng.uniform(-60, 60, (3000, 3))
    m[:, 3] = rng.uniform(0, 10, 3000)
    T = np.eye(4)
    T[:3, :3] = np.linalg.qr(rng.normal(size=(3, 3)))[0]
    T[:3, 3] = rng.normal(size=3)
    out = po.reanchor(m, T, 8.0, 3.0, 50.0, False, 0.3)
    keep = (m[:, 3].astype(np.float64) + 3.0 >= 8.0) & (np.linalg.norm(m[:, :3].astype(np.float64), axis=1) <= 50.0 + 1e-4)
    ref = (m[keep, :3].astype(np.float64) @ T[:3, :3].T + T[:3, 3]).astype(np.float32)
    assert abs(len(out) - keep.sum()) <= 1  # a point within fp32 rounding of the distance gate may flip
    if len(out) == keep.sum():
        assert np.abs(out[:, :3] - ref).max() <= 1e-5 and np.array_equal(out[:, 3], m[keep, 3])


def test_chi2_quantile_table_matches_scipy(po):
    """UpdaterStatistics.cpp:31-33: chi_squared_table[i] = quantile(chi_squared(i), 0.95); the oracle computes it itself."""
    from scipy.stats import chi2
    for dof in range(1, 40):
        assert po.chi2_quantile95(dof) == pytest.approx(float(chi2.ppf(0.95, dof)), rel=1e-13)


@pytest.mark.parametrize("k", [3, 4, 5, 8, 10])
def test_oracle_chi2_gate_follows_the_projected_row_count(po, workload, k):
    """get_chi2 reads chi_squared_table[res.rows()] with res.rows() = k - 2 after the projection (UpdaterStatistics.cpp:110-111)."""
    import copy
    from scipy.stats import chi2
    w = workload("G-small", k=k)
    op = w["op"]
    sc = po.range_filter(w["raw"], w["t_rel"], op.max_range, op.min_range)
    sm = po.transform_f32(sc, w["T_LtoM"])
    r0 = po.scan_linsys(w["consts"], w["P"], w["map"], sc, sm)
    tested = (r0["rows"]["status"] == 0) | (r0["rows"]["status"] == 4)
    q = float(chi2.ppf(0.95, k - 2))
    c2 = copy.copy(w["consts"])
    v = np.sort(r0["rows"]["chi2"][tested])
    c2.chi2_mult = float(0.5 * (v[len(v) // 2] + v[len(v) // 2 + 1])) / q      # between two points: nobody sits on the gate
    r1 = po.scan_linsys(c2, w["P"], w["map"], sc, sm)
    st = r1["rows"]["status"]
    assert np.array_equal(st[tested] == 0, r1["rows"]["chi2"][tested] < c2.chi2_mult * q)
    assert 0.3 * tested.sum() < (st == 4).sum() < 0.7 * tested.sum()
    assert r1["total_res"] == (k - 2) * (st == 0).sum()


def test_oracle_plane_gate_on_rank_deficient_neighbours(po):
    """LidarHelper.cpp:305-312: collinear / coplanar-with-the-origin neighbours give sigma_max / sigma_min = huge or inf."""
    t = np.arange(5, dtype=np.float32)[:, None]
    for nbp in (t * np.float32([0.3, 0.3, 0.0]) + np.float32([0.3, 0.3, 0.0]), t * np.float32([0.25, 0.0, 0.0]) + np.float32([3.0, 2.0, 0.5]),
                np.float32([[2, 0, 0], [2.3, 0, 0], [2, 0.3, 0], [2.3, 0.3, 0], [2.15, 0.15, 0]])):
        mp = np.zeros((5, 4), np.float32)
        mp[:, :3] = nbp
        ok, plane, cond = po.compute_plane(mp)
        assert not ok and cond > 1e12


def _newton_eig_minmax(G):
    """numpy restatement of rows.cuh sym3_eig_minmax: roots of 4c^3 - 3c = r by Newton instead of cos(acos(r)/3)."""
    a00, a01, a02, a11, a12, a22 = G[0, 0], G[0, 1], G[0, 2], G[1, 1], G[1, 2], G[2, 2]
    p1 = a01 * a01 + a02 * a02 + a12 * a12
    q = (a00 + a11 + a22) / 3.0
    b00, b11, b22 = a00 - q, a11 - q, a22 - q
    p2 = b00 * b00 + b11 * b11 + b22 * b22 + 2.0 * p1
    if not p2 > 0.0:
        return q, q
    p = np.sqrt(p2 / 6.0)
    det = b00 * (b11 * b22 - a12 * a12) - a01 * (a01 * b22 - a12 * a02) + a02 * (a01 * a12 - b11 * a02)
    r = min(1.0, max(-1.0, 0.5 * det / p**3))
    c = np.float32(0.5) + np.sqrt((np.float32(1.0) + np.float32(r)) / np.float32(6.0))
    rf = np.float32(r)
    for _ in range(3):
        c2 = c * c
        c = np.float32(c - ((np.float32(4) * c2 - np.float32(3)) * c - rf) / (np.float32(12) * c2 - np.float32(3) + np.float32(1e-30)))
    c = max(0.5, float(c))
    for _ in range(2):
        fp = 12.0 * c * c - 3.0
        if fp > 1e-300:
            c = c - ((4.0 * c * c - 3.0) * c - r) / fp
    c = min(1.0, max(0.5, c))
    return q - p * (np.sqrt(3.0) * np.sqrt(max(0.0, 1.0 - c * c)) + c), q + 2.0 * p * c


@pytest.mark.parametrize("k", [3, 5, 10])
def test_gram_route_matches_qr_route(po, k):
    """The r2 device route of compute_plane (rows.cuh point_row): G = A^T A, condition gate from the extreme eigenvalues of G found by
    Newton on 4c^3 - 3c = r, abc = -d G^-1 A^T 1 by a 3x3 Cholesky - against the oracle's SVD + column-pivoted QR restatement of
    LidarHelper.cpp:291-328 on neighbour sets that pass the reference's gate (cond <= plane_max_condi)."""
    rng = np.random.default_rng(100 + k)
    n = 0
    for trial in range(400):
        off = rng.uniform(-6, 6, 3)
        nb = plane_neighbours(rng, k, noise=rng.choice([0.0, 0.005, 0.02, 0.1]), offset=off)
        max_condi = 100.0
        ok, plane, cond = po.compute_plane(nb, max_condi, 1e9)
        A = nb[:, :3].astype(np.float64)
        G = A.T @ A
        lmin, lmax = _newton_eig_minmax(G)
        ev = np.linalg.eigvalsh(G)
        assert lmax == pytest.approx(ev[-1], rel=1e-13)
        assert lmin == pytest.approx(ev[0], rel=1e-9 * cond * cond, abs=1e-13 * ev[-1])
        gate = lmin > 0.0 and not (lmax > max_condi**2 * lmin)
        if abs(cond - max_condi) > 1e-6 * max_condi:
            assert gate == (cond <= max_condi)
        if not (gate and ok):
            continue
        d = np.linalg.norm(A, axis=1).mean()
        L = np.linalg.cholesky(G)
        abc = -d * np.linalg.solve(L.T, np.linalg.solve(L, A.sum(0)))
        assert np.abs(abc - plane[:3]).max() <= 1e-15 * cond**2 * np.abs(plane[:3]).max() + 1e-300
        assert np.abs(abc - plane[:3]).max() <= 3e-11 * np.abs(plane[:3]).max()
        n += 1
    assert n > 25
