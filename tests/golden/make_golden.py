"""Generates the committed golden fixtures from the REFERENCE ITSELF, run in the build container:

  * nabo_scan3d.npz / nabo_scan3d_large.npz — libnabo's own kNN validation clouds
    (thirdparty/libnabo/tests/data/scan.3d.txt, scan.3d.large.txt; used by tests/knnvalidate.cpp:74-238 with K=10)
    plus the k-nearest neighbours returned by the reference's ikd-Tree (mins/src/update/lidar/ikd_Tree.cpp compiled
    unmodified -> oracle/_ref/libikd_ref.so) for seeded queries: neighbour coordinates and squared distances.
  * ikd_add_points.npz — KD_TREE::Build + Add_Points(downsample=true) on a seeded cloud: the resulting live point set.

/root/reference is only readable here, so the outputs are committed; tests never read /root/reference.
Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import pyoracle as po  # noqa: E402

REF = "/root/reference/thirdparty/libnabo/tests/data"


def nabo(name, fn, n_query, seed, max_pts=None):
    cloud = np.loadtxt(os.path.join(REF, fn), dtype=np.float64)[:, :3].astype(np.float32)
    rng = np.random.default_rng(seed)
    if max_pts and len(cloud) > max_pts:
        cloud = cloud[np.sort(rng.choice(len(cloud), max_pts, replace=False))]
    lo, hi = cloud.min(0), cloud.max(0)
    q1 = rng.uniform(lo, hi, (n_query // 2, 3))                       # knnvalidate "random" queries: inside the bounds
    sel = rng.choice(len(cloud), n_query - n_query // 2, replace=False)
    q2 = cloud[sel] + rng.normal(0, 0.05, (len(sel), 3))               # near the surface, like a registered scan
    q = np.concatenate([q1, q2]).astype(np.float32)
    xyzi = np.zeros((len(cloud), 4), np.float32)
    xyzi[:, :3] = cloud
    xyzi[:, 3] = np.arange(len(cloud), dtype=np.float32)               # intensity = original index (exact below 2^24)
    q4 = np.zeros((len(q), 4), np.float32)
    q4[:, :3] = q
    tree = po.RefIkd(0.3)
    tree.build(xyzi)
    nb, d2, found = tree.knn(q4, 10, threads=4)
    assert (found == 10).all()
    np.savez_compressed(os.path.join(HERE, name), cloud=cloud, queries=q, ref_nb=nb, ref_d2=d2)
    print(name, cloud.shape, q.shape)


def ikd_add(seed=3):
    rng = np.random.default_rng(seed)
    # a wavy surface patch sampled densely (several points per 0.3 m voxel) so Build keeps duplicates per voxel
    n0, n1 = 6000, 5000
    def surf(n):
        xy = rng.uniform(-6, 6, (n, 2))
        z = 0.3 * np.sin(xy[:, 0]) + 0.2 * np.cos(1.7 * xy[:, 1]) + rng.normal(0, 0.02, n)
        return np.concatenate([xy, z[:, None]], 1)
    base = np.zeros((n0, 4), np.float32)
    base[:, :3] = surf(n0)
    base[:, 3] = rng.uniform(0, 1, n0)
    add = np.zeros((n1, 4), np.float32)
    add[:, :3] = surf(n1) + np.array([3.0, 0, 0])                      # half overlapping, half new ground
    add[:, 3] = rng.uniform(1, 2, n1)
    tree = po.RefIkd(0.3)
    tree.build(base)
    tree.add(add, True)
    out = tree.flatten()
    out = out[np.lexsort((out[:, 3], out[:, 2], out[:, 1], out[:, 0]))]
    np.savez_compressed(os.path.join(HERE, "ikd_add_points"), base=base, add=add, ds=np.float32(0.3), result_sorted=out)
    print("ikd_add_points", base.shape, add.shape, out.shape)


if __name__ == "__main__":
    nabo("nabo_scan3d", "scan.3d.txt", 2000, 1)
    nabo("nabo_scan3d_large", "scan.3d.large.txt", 1000, 2, max_pts=30000)
    ikd_add()
