"""TEST INFRASTRUCTURE ONLY — CPU restatement (numpy, fp32 where the reference is `PointMatcher<float>`) of the ICP block of
LidarHelper::transform_to_map (mins/src/update/lidar/LidarHelper.cpp:156-187), i.e. the libpointmatcher chain
`icp.setDefault()` + {MaxDist, SamplingSurfaceNormal} reference filters (thirdparty/libpointmatcher/pointmatcher/ICP.cpp:
100-113, 265-449). Imported only by tests/.

Parity statement: **unpinned against the real library** — libpointmatcher needs Boost/Eigen/libnabo, none of which exist in
this image, so it cannot be compiled here; the reference's own tests hold no golden vector for this chain; and the chain is
non-deterministic in the reference (RandomSampling seeds `std::minstd_rand` from `std::random_device`, SamplingSurfaceNormal
draws from `std::rand`). What follows each function is the file:line it restates. The 1-NN search itself IS pinned: it is
exact nearest neighbour (libnabo, epsilon = 0), checked against the reference's libnabo fixtures elsewhere in tests/.
"""
import numpy as np
from scipy.spatial import cKDTree

F = np.float32


def random_sampling(n, prob, rng):
    """RandomSampling.cpp:92-109: keep point i iff u_i < prob, stop once more than floor(n * prob) were kept."""
    nb_out = int(F(n) * F(prob))
    u = rng.random(n).astype(F)
    keep = []
    for i in range(n):
        if len(keep) > nb_out:
            break
        if u[i] < prob:
            keep.append(i)
    return np.array(keep, np.int64)


def max_dist_filter(pts, max_dist):
    """MaxDist.cpp (dim = -1): keep points with ||p|| < maxDist, fp32 norm."""
    nrm = np.sqrt((pts.astype(F) ** 2).sum(1, dtype=F))
    return np.where(nrm < F(abs(max_dist)))[0]


def sampling_surface_normal(pts, rng, ratio=0.5, knn=7):
    """SamplingSurfaceNormal.cpp:170-296 (samplingMethod 0, keepNormals): recursive median split along the largest extent of
    the inherited box until <= knn points, PCA normal of the bin (eigenvector of the smallest eigenvalue of the bin's
    covariance), every point of the bin kept with probability `ratio` carrying that normal. Returns (kept indices sorted,
    normals of the kept points)."""
    pts = np.ascontiguousarray(pts, F)
    n = len(pts)
    idx = np.arange(n)
    keep, normals = [], []
    stack = [(0, n, pts.min(0).copy(), pts.max(0).copy())]
    while stack:
        first, last, mn, mx = stack.pop()
        count = last - first
        if count <= knn:
            b = idx[first:last]
            d = pts[b]
            mean = d.sum(0, dtype=F) / F(count)
            NN = d - mean
            C = (NN.T @ NN) / F(count)
            if np.linalg.matrix_rank(C.astype(np.float64)) + 1 < 3:  # :247 fullPivHouseholderQr().rank() + 1 >= featDim - 1
                continue
            w, v = np.linalg.eigh(C.astype(np.float64))
            nrm = v[:, 0].astype(F)  # utils.h computeNormal: eigenvector of the smallest eigenvalue
            r = rng.random(count).astype(F)  # :281 (float)std::rand() / RAND_MAX
            for j in range(count):
                if r[j] < ratio:
                    keep.append(b[j])
                    normals.append(nrm)
            continue
        cut = int(np.argmax(mx - mn))                      # :184
        right = count // 2
        left = count - right
        seg = idx[first:last]
        order = np.argpartition(pts[seg, cut], left) if left < count else np.arange(count)  # :192-197 nth_element
        idx[first:last] = seg[order]
        cutval = pts[idx[first + left], cut]               # :200-201
        lmx = mx.copy()
        lmx[cut] = cutval
        rmn = mn.copy()
        rmn[cut] = cutval
        # recursion order of the reference is left then right (std::rand sequence); a stack pops right first, so push
        # right before left
        stack.append((first + left, last, rmn, mx))
        stack.append((first, first + left, mn, lmx))
    keep = np.array(keep, np.int64)
    normals = np.array(normals, F).reshape(-1, 3)
    o = np.argsort(keep, kind="stable")                    # :139 std::sort(indicesToKeep)
    return keep[o], normals[o]


def point_to_plane_step(read_c, ref_c, nrm, w):
    """PointToPlane.cpp:172-283 for 3-D, 6 DoF: F = [p x n ; n], A = wF F^T, b = -(wF ((p - r).n)), x = A.llt().solve(b),
    rotation AngleAxis(|x[:3]|, x[:3]/|x[:3]|), translation x[3:]. fp32 like PointMatcher<float>."""
    p, r, n = read_c.astype(F), ref_c.astype(F), nrm.astype(F)
    Fm = np.concatenate([np.cross(p, n), n], 1).astype(F)          # (m, 6)
    wF = Fm * w[:, None].astype(F)
    A = (wF.T @ Fm).astype(F)
    dot = ((p - r) * n).sum(1, dtype=F)
    b = -(wF.T @ dot).astype(F)
    x = np.linalg.solve(A.astype(np.float64), b.astype(np.float64))
    ang = np.linalg.norm(x[:3])
    T = np.eye(4)
    if ang > 0:
        u = x[:3] / ang
        K = np.array([[0, -u[2], u[1]], [u[2], 0, -u[0]], [-u[1], u[0], 0]])
        T[:3, :3] = np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * (K @ K)
    T[:3, 3] = x[3:]
    return T.astype(F)


def knn_surface_normal(pts, rng, ratio=0.5, knn=7):
    """The DEVICE's variant of the reference filter (mins_b200/csrc/icp.cuh k_icp_ref_select + k_icp_normals): every point kept with
    probability `ratio`, normal = PCA of its `knn` nearest points of the unfiltered cloud (itself included), rank-deficient
    neighbourhoods dropped. Same support size as the reference's bins, no sequential recursion. Returns (kept indices, normals)."""
    pts = np.ascontiguousarray(pts, F)
    keep = np.where(rng.random(len(pts)) < ratio)[0]
    _, nn = cKDTree(pts.astype(np.float64)).query(pts[keep].astype(np.float64), k=knn)
    d = pts[nn].astype(np.float64)                                   # (m, knn, 3)
    d = d - d.mean(1, keepdims=True)
    C = np.einsum("mki,mkj->mij", d, d) / knn
    w, v = np.linalg.eigh(C)
    ok = w[:, 1] > 1e-12 * np.maximum(w[:, 2], 1e-300)               # rank >= 2
    return keep[ok], v[ok, :, 0].astype(F)


def icp(reading_xyz, map_xyz, T_init, map_icp_dist, seed=0, read_prob=0.75, ref_ratio=0.5, knn=7, trim=0.85, max_iter=40,
        min_rot=1e-3, min_trans=1e-3, smooth=3, normals="bins", max_match_dist=None, stats=None):
    # normals = "bins": the reference's median-split bins; "knn": the device's 7-NN variant. max_match_dist = None: the reference's
    # unbounded matcher (MatchersImpl.h:83-86); a number: the device's bound (reading points without a reference point within it
    # are dropped before the trimmed filter). stats: optional dict that receives per-iteration match statistics.
    """PointMatcher<float>::ICP::compute (ICP.cpp:265-449) + the success rule of LidarHelper.cpp:176-184.
    Returns dict(T_LtoM, T_icp, success, iterations, max_iter_reached, n_reading, n_reference)."""
    rng = np.random.default_rng(seed)
    T_init = np.asarray(T_init, F).reshape(4, 4)
    ref = np.ascontiguousarray(map_xyz, F)[:, :3]
    ref = ref[max_dist_filter(ref, map_icp_dist)]
    kidx, nrm = (sampling_surface_normal if normals == "bins" else knn_surface_normal)(ref, rng, ref_ratio, knn)
    ref = ref[kidx]
    mean = (ref.sum(0, dtype=F) / F(len(ref))).astype(F)               # ICP.cpp:292-294
    ref_c = ref - mean                                                 # :300
    tree = cKDTree(ref_c.astype(np.float64))
    rd = np.ascontiguousarray(reading_xyz, F)[:, :3]
    rd = rd[random_sampling(len(rd), read_prob, rng)]
    T_rm = np.eye(4, dtype=F)
    T_rm[:3, 3] = mean
    T_refMean_dataIn = (np.linalg.inv(T_rm.astype(np.float64)) @ T_init.astype(np.float64)).astype(F)   # :351
    reading = (rd @ T_refMean_dataIn[:3, :3].T + T_refMean_dataIn[:3, 3]).astype(F)
    T_iter = np.eye(4, dtype=F)
    hist = [T_iter.copy()]
    it, max_reached = 0, False
    while True:
        step = (reading @ T_iter[:3, :3].T + T_iter[:3, 3]).astype(F)                                 # :387
        d, j = tree.query(step.astype(np.float64), k=1)                                              # :391 exact 1-NN
        d2 = (d * d).astype(F)
        if stats is not None:
            stats.setdefault("max_match", []).append(float(d.max()))
            stats.setdefault("beyond_2m", []).append(int((d > 2.0).sum()))
            stats.setdefault("trim_limit", []).append(float(np.sqrt(np.partition(d2, int(F(len(d2)) * F(trim)))[int(F(len(d2)) * F(trim))])))
        if max_match_dist is not None:
            inb = d <= max_match_dist
            step, j, d2 = step[inb], j[inb], d2[inb]
        q = int(F(len(d2)) * F(trim))                                                                # Matches.cpp:85-86
        limit = np.partition(d2, q)[q] if trim < 1.0 else d2.max()
        w = (d2 <= limit).astype(F)                                                                  # OutlierFiltersImpl.cpp:146
        m = w != 0
        T_iter = (point_to_plane_step(step[m], ref_c[j[m]], nrm[j[m]], w[m]).astype(np.float64) @ T_iter.astype(np.float64)).astype(F)
        it += 1
        if it >= max_iter:                                                                           # Counter checker
            max_reached = True
            break
        hist.append(T_iter.copy())
        if len(hist) > smooth:                                                                       # Differential checker
            dr = dt = 0.0
            for i in range(len(hist) - 1, len(hist) - 1 - smooth, -1):
                Rrel = hist[i][:3, :3].astype(np.float64) @ hist[i - 1][:3, :3].astype(np.float64).T
                dr += abs(np.arccos(np.clip((np.trace(Rrel) - 1) / 2, -1, 1)))
                dt += np.linalg.norm(hist[i][:3, 3] - hist[i - 1][:3, 3])
            if dr / smooth < min_rot and dt / smooth < min_trans:
                break
    T_icp = (T_rm.astype(np.float64) @ T_iter.astype(np.float64) @ T_refMean_dataIn.astype(np.float64)).astype(F)   # :447
    T_diff = T_icp.astype(np.float64) @ np.linalg.inv(T_init.astype(np.float64))                     # LidarHelper.cpp:176
    ok = (not max_reached) and not (np.linalg.norm(T_diff[:3, 3]) > 1)
    return dict(T_LtoM=T_icp if ok else T_init, T_icp=T_icp, success=ok, iterations=it, max_iter_reached=max_reached,
                n_reading=len(rd), n_reference=len(ref))
