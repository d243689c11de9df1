"""CPU check of the search STRUCTURE the device uses instead of the ikd-Tree (oracle/grid_walk.py restates map.cuh's grid and
update.cuh's cube walk): the visiting rule + completeness test return exactly the brute-force k nearest neighbours under the
total order (d2 fp32 non-contracted, map index), with get_neighbors' gate (LidarHelper.cpp:280-285), also where the GPU workloads
rarely go: queries far outside the grid, one-cell grids, fewer points than k, duplicate points, a cell edge doubled by the cap."""
import numpy as np
import pytest

from oracle import grid_walk as gw


def surface_cloud(rng, n, extent=12.0):
    """points on a few planes and a blob: the density contrasts of a voxel-thinned map"""
    a = rng.uniform(-extent, extent, (n // 2, 3)); a[:, 2] = -1.5 + 0.01 * rng.standard_normal(len(a))
    b = rng.uniform(-extent, extent, (n // 4, 3)); b[:, 0] = extent + 0.01 * rng.standard_normal(len(b))
    c = rng.normal(0, 0.7, (n - len(a) - len(b), 3)) + [3.0, -2.0, 0.5]
    return np.concatenate([a, b, c]).astype(np.float32)


def check(g, pts, queries, k, max_d2):
    n_ok = 0
    for q in queries:
        st = {}
        i0, d0, ok0 = gw.knn_walk(g, q, k, max_d2, st)
        i1, d1, ok1 = gw.knn_brute(pts, q, k, max_d2)
        assert ok0 == ok1 and np.array_equal(i0, i1) and np.array_equal(d0, d1), (q, i0, i1)
        n_ok += ok0
    return n_ok


@pytest.mark.parametrize("k,max_d2,cell0", [(5, 1.0, 0.5), (1, 4.0, 0.5), (10, 10.0, 0.5), (5, 1.0, 2.0), (3, 0.3, 0.25)])
def test_cube_walk_equals_brute_force(k, max_d2, cell0):
    rng = np.random.default_rng(k * 100 + int(cell0 * 8))
    pts = surface_cloud(rng, 6000)
    g = gw.Grid(pts, cell0=cell0)
    inside = pts[rng.choice(len(pts), 120)] + rng.normal(0, 0.15, (120, 3)).astype(np.float32)
    sparse = rng.uniform(-14, 14, (80, 3)).astype(np.float32)                 # mostly empty space inside the box
    outside = (rng.uniform(-1, 1, (40, 3)) * [40, 40, 10] + [0, 0, 20]).astype(np.float32)   # far outside the grid
    n_ok = check(g, pts, np.concatenate([inside, sparse, outside]), k, max_d2)
    assert n_ok >= 100                                                         # the near-surface queries do find their neighbours


def test_cell_edge_doubles_under_the_cap_and_results_do_not_change():
    rng = np.random.default_rng(3)
    pts = surface_cloud(rng, 3000)
    g_small = gw.Grid(pts, cell0=0.5, max_cells=4000)                          # 48 x 48 x ~10 cells would be 23 000
    g_ref = gw.Grid(pts, cell0=0.5)
    assert g_small.cell in (1.0, 2.0, 4.0) and g_ref.cell == 0.5 and int(g_small.dims.prod()) <= 4000
    q = pts[rng.choice(len(pts), 60)] + rng.normal(0, 0.2, (60, 3)).astype(np.float32)
    for g in (g_small, g_ref):
        check(g, pts, q, 5, 1.0)


def test_degenerate_maps():
    rng = np.random.default_rng(5)
    # fewer points than k; all points in one cell; duplicates (ties are broken by the map index)
    few = np.float32([[0.1, 0.1, 0.1], [0.2, 0.1, 0.1], [0.1, 0.3, 0.1]])
    g = gw.Grid(few)
    for q in np.float32([[0.1, 0.1, 0.1], [0.4, 0.4, 0.4], [5, 5, 5]]):
        i, d, ok = gw.knn_walk(g, q, 5, 1.0)
        assert not ok and np.array_equal(i, gw.knn_brute(few, q, 5, 1.0)[0])
    dup = np.repeat(surface_cloud(rng, 200), 3, axis=0)
    dup = dup[rng.permutation(len(dup))]
    g = gw.Grid(dup)
    q = dup[rng.choice(len(dup), 40)]
    check(g, dup, q, 5, 1.0)
    i, d, ok = gw.knn_walk(g, q[0], 5, 1.0)
    assert d[0] == d[1] == d[2] == 0.0 and i[0] < i[1] < i[2]                  # three copies of the query point, index order


def test_the_gate_compares_the_squared_distance_with_the_unsquared_option():
    """map_ngbr_max_d = 1.0 m is used as a SQUARED distance by get_neighbors (LidarHelper.cpp:284); with 0.25 the gate is at 0.5 m"""
    pts = np.float32([[0, 0, 0], [0.1, 0, 0], [0.2, 0, 0], [0.3, 0, 0], [0.6, 0, 0], [3, 3, 3]])
    g = gw.Grid(pts)
    i, d, ok = gw.knn_walk(g, np.float32([0, 0, 0]), 5, 0.25)
    assert not ok and list(i) == [0, 1, 2, 3, -1] and np.isinf(d[4])           # the 5th neighbour is 0.6 m away: 0.36 > 0.25
    i, d, ok = gw.knn_walk(g, np.float32([0, 0, 0]), 5, 0.37)
    assert ok and list(i) == [0, 1, 2, 3, 4]
