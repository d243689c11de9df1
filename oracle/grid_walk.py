"""TEST INFRASTRUCTURE ONLY - a numpy restatement of the search STRUCTURE that replaces the ikd-Tree on the device: the dense grid
of mins_b200/csrc/map.cuh (grid_setup_dev, cell_of) and the exact cube walk of mins_b200/csrc/update.cuh (knn_walk, knn_walk_staged,
knn_walk_lane: three schedules of one visiting rule). It exists so that the exactness argument of DESIGN.md 3.1 can be run on a
CPU against brute force, including the corners the GPU parity workloads rarely reach: queries far outside the grid, grids of one
cell, fewer map points than k, duplicate points (ties by index), the level cap R and the reference's squared-vs-unsquared gate
(LidarHelper.cpp:284). It says nothing about the CUDA code itself: that is what tests/test_gpu_parity*.py are for.
"""
import numpy as np

F = np.float32


class Grid:
    """grid_setup_dev: origin one cell below the floor of the bounding box, dims cover the box plus padding, the cell edge doubles
    until dimx * dimy * dimz <= max_cells; cell_of clamps a point's cell to the grid."""

    def __init__(self, pts, cell0=0.5, max_cells=1 << 24):
        p = np.ascontiguousarray(pts, F)[:, :3]
        self.pts = p
        mn, mx = p.min(0).astype(np.float64), p.max(0).astype(np.float64)
        cell = float(cell0)
        while True:
            o = np.floor(mn / cell) * cell - cell
            d = np.floor((mx - o) / cell).astype(np.int64) + 2
            if int(d[0]) * int(d[1]) * int(d[2]) <= max_cells:
                break
            cell *= 2.0
        self.o, self.dims, self.cell, self.inv_cell = o, d.astype(np.int64), cell, 1.0 / cell
        c = np.floor((p.astype(np.float64) - o) * self.inv_cell).astype(np.int64)
        c = np.minimum(np.maximum(c, 0), self.dims - 1)
        cid = (c[:, 2] * self.dims[1] + c[:, 1]) * self.dims[0] + c[:, 0]
        self.order = np.argsort(cid, kind="stable")            # cell_pts: the points regrouped by cell (w = map index)
        self.cell_start = np.zeros(int(self.dims.prod()) + 1, np.int64)
        np.add.at(self.cell_start, cid + 1, 1)
        self.cell_start = np.cumsum(self.cell_start)

    def cell_coord(self, q):
        return np.floor((np.asarray(q, np.float64) - self.o) * self.inv_cell).astype(np.int64)


def dist2_ref(a, b):
    """KD_TREE::calc_dist (ikd_Tree.cpp:1406-1410) without contraction: ((dx dx) + (dy dy)) + (dz dz) in fp32"""
    d = (np.asarray(a, F) - np.asarray(b, F)).astype(F)
    return F(F(F(d[..., 0] * d[..., 0]) + F(d[..., 1] * d[..., 1])) + F(d[..., 2] * d[..., 2]))


def knn_walk(g, q, k, max_d2, stats=None):
    """One query. Returns (idx[k] (-1 = none within the gate), d2[k], ok) like ml_knn: candidates beyond max_d2 are not reported,
    ok = k neighbours found and the k-th squared distance <= max_d2 (get_neighbors, LidarHelper.cpp:280-285)."""
    q = np.asarray(q, F)
    lim = 1 << 19
    c = np.clip(g.cell_coord(q), -lim, lim)
    f = (q.astype(np.float64) - g.o) * g.inv_cell - c
    fmn = max(0.0, float(np.minimum(f, 1.0 - f).min()))
    dx, dy, dz = (int(v) for v in g.dims)
    cx, cy, cz = (int(v) for v in c)
    R = max(1, int(np.ceil(np.sqrt(np.float64(F(max_d2))) * (1.0 + 1e-5) * g.inv_cell)))
    s_lo = max(1, max(-cx, cx - (dx - 1), 0), max(-cy, cy - (dy - 1), 0), max(-cz, cz - (dz - 1), 0))
    s_hi = min(R, max(cx, dx - 1 - cx, cy, dy - 1 - cy, cz, dz - 1 - cz))
    best = []                                                  # (d2 bits as float, map index), kept sorted, at most k
    visited = 0
    for s in range(s_lo, s_hi + 1):
        y0, y1 = max(cy - s, 0), min(cy + s, dy - 1)
        z0, z1 = max(cz - s, 0), min(cz + s, dz - 1)
        for y in range(y0, y1 + 1):
            for z in range(z0, z1 + 1):
                base = (z * dy + y) * dx
                outer = s == s_lo or abs(y - cy) == s or abs(z - cz) == s
                if outer:
                    x0, x1 = max(cx - s, 0), min(cx + s, dx - 1)
                    rng = [(g.cell_start[base + x0], g.cell_start[base + x1 + 1])] if x0 <= x1 else []
                else:
                    rng = [(g.cell_start[base + x], g.cell_start[base + x + 1]) for x in (cx - s, cx + s) if 0 <= x < dx]
                for b, e in rng:
                    for j in range(int(b), int(e)):
                        idx = int(g.order[j])
                        d2 = float(dist2_ref(g.pts[idx], q))
                        visited += 1
                        if len(best) < k or (d2, idx) < best[-1]:
                            best.append((d2, idx))
                            best.sort()
                            del best[k:]
        gs = (fmn + s) * g.cell
        if len(best) == k and best[-1][0] < gs * gs * (1.0 - 1e-6):
            break
    if stats is not None:
        stats["visited"] = visited
    found = len(best) == k
    ok = found and not (np.float64(F(best[-1][0])) > np.float64(F(max_d2)))
    idx = np.full(k, -1, np.int32)
    d2 = np.full(k, np.inf, F)
    for j, (d, i) in enumerate(best):
        if d <= F(max_d2):
            idx[j], d2[j] = i, d
    return idx, d2, bool(ok)


def knn_brute(pts, q, k, max_d2):
    """the same contract by exhaustive search under the total order (d2, index)"""
    p = np.ascontiguousarray(pts, F)[:, :3]
    d2 = dist2_ref(p, np.asarray(q, F)[None, :])
    o = np.lexsort((np.arange(len(p)), d2))[:k]
    found = len(o) == k
    ok = found and not (np.float64(d2[o[-1]]) > np.float64(F(max_d2)))
    idx = np.full(k, -1, np.int32)
    out = np.full(k, np.inf, F)
    for j, i in enumerate(o):
        if d2[i] <= F(max_d2):
            idx[j], out[j] = i, d2[i]
    return idx, out, bool(ok)
