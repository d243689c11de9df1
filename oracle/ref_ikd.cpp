// TEST INFRASTRUCTURE ONLY — never linked into the product library.
// C wrapper around the reference's own, unmodified ikd-Tree (mins/src/update/lidar/ikd_Tree.{h,cpp}),
// which oracle/Makefile compiles from /root/reference in place into oracle/_ref/libikd_ref.so.
// It exposes exactly the calls the LiDAR path makes:
//   KD_TREE::Build            <- LidarHelper::init_map_local      (LidarHelper.cpp:103-113)
//   KD_TREE::Nearest_Search   <- LidarHelper::get_neighbors       (LidarHelper.cpp:268-289)
//   KD_TREE::Add_Points       <- LidarHelper::register_scan       (LidarHelper.cpp:196-207)
//   KD_TREE::flatten          <- propagate_map_to_newest_clone    (LidarHelper.cpp:213)
// Points cross the boundary as float[4] = {x, y, z, intensity}.
#include "ikd_Tree.h"
#include <thread>
#include <vector>

typedef pcl::PointXYZI P;
typedef KD_TREE<P> Tree;
typedef Tree::PointVector PV;

static PV to_pv(const float *xyzi, int n) {
  PV v(n);
  for (int i = 0; i < n; i++) {
    v[i].x = xyzi[4 * i + 0];
    v[i].y = xyzi[4 * i + 1];
    v[i].z = xyzi[4 * i + 2];
    v[i].intensity = xyzi[4 * i + 3];
  }
  return v;
}

extern "C" {
void *ref_ikd_create(float downsample_size) {
  Tree *t = new Tree(); // ~80 MB object (operation queue), must live on the heap
  t->set_downsample_param(downsample_size);
  return t;
}
void ref_ikd_destroy(void *h) { delete (Tree *)h; }
void ref_ikd_build(void *h, const float *xyzi, int n) { ((Tree *)h)->Build(to_pv(xyzi, n)); }
int ref_ikd_size(void *h) { return ((Tree *)h)->validnum(); }
int ref_ikd_add(void *h, const float *xyzi, int n, int downsample) {
  PV v = to_pv(xyzi, n);
  return ((Tree *)h)->Add_Points(v, downsample != 0);
}
int ref_ikd_flatten(void *h, float *out, int cap) {
  Tree *t = (Tree *)h;
  PV v;
  t->flatten(t->Root_Node, v, NOT_RECORD);
  int n = (int)v.size() < cap ? (int)v.size() : cap;
  for (int i = 0; i < n; i++) {
    out[4 * i + 0] = v[i].x;
    out[4 * i + 1] = v[i].y;
    out[4 * i + 2] = v[i].z;
    out[4 * i + 3] = v[i].intensity;
  }
  return (int)v.size();
}
// k-NN of nq queries (float[4] each, w ignored). Outputs neighbour coordinates (float[4]) and squared
// distances, ascending, exactly as Nearest_Search returns them; found[i] = number returned (<= k).
// Threading mirrors UpdaterLidar.cpp:269-277 (static chunks, last thread takes the remainder).
void ref_ikd_knn(void *h, const float *q, int nq, int k, float *nb_xyzi, float *d2, int *found, int threads) {
  Tree *t = (Tree *)h;
  if (threads < 1)
    threads = 1;
  int chunk = nq / threads;
  std::vector<std::thread> th;
  for (int ti = 0; ti < threads; ti++) {
    int s = ti * chunk, e = (ti == threads - 1) ? nq : (ti + 1) * chunk;
    th.emplace_back([=] {
      PV nb;
      std::vector<float> dd;
      for (int i = s; i < e; i++) {
        P p;
        p.x = q[4 * i];
        p.y = q[4 * i + 1];
        p.z = q[4 * i + 2];
        t->Nearest_Search(p, k, nb, dd);
        found[i] = (int)nb.size();
        for (int j = 0; j < k; j++) {
          bool ok = j < (int)nb.size();
          nb_xyzi[(size_t)(i * (size_t)k + j) * 4 + 0] = ok ? nb[j].x : 0.f;
          nb_xyzi[(size_t)(i * (size_t)k + j) * 4 + 1] = ok ? nb[j].y : 0.f;
          nb_xyzi[(size_t)(i * (size_t)k + j) * 4 + 2] = ok ? nb[j].z : 0.f;
          nb_xyzi[(size_t)(i * (size_t)k + j) * 4 + 3] = ok ? nb[j].intensity : 0.f;
          d2[(size_t)i * k + j] = ok ? dd[j] : INFINITY;
        }
      }
    });
  }
  for (auto &x : th)
    x.join();
}
}

// ---------------------------------------------------------------------------------------------------------
// Whole CPU LiDAR update of one scan, as SURVEY.md §8(d) defines the timed CPU path: the reference's own
// KD_TREE::Nearest_Search inside the static-chunk thread loop of UpdaterLidar.cpp:269-277, followed by the restated
// rows / stacking / Givens compression (oracle/lidar_oracle.cpp). Declared here, resolved from liboracle.so.
// ---------------------------------------------------------------------------------------------------------
#include "../include/mins_lidar.h"
extern "C" int orc_range_filter(const float *xyz, int stride, int n, float t_rel, double max_range, double min_range, float *out_xyzi);
extern "C" void orc_transform_f32(const float *in_xyzi, int n, const float *T, float *out_xyzi);
extern "C" int orc_scan_linsys_nb(const ml_scan_consts *c, const double *P, const float *scan_L, int N, const float *nb, const uint8_t *nb_ok,
                                  double *HtH, double *Htr, ml_counters *cnt, double *Hc, double *rc, int *n_c, int threads);

extern "C" int ref_lidar_update(void *tree, const float *raw, int stride, int n, float t_rel, double rmax, double rmin, const float *T_LtoM,
                                const ml_scan_consts *c, const double *P, double *Hc, double *rc, int *n_c, ml_counters *cnt, int threads,
                                double *HtH, double *Htr) {
  std::vector<float> sc(4 * (size_t)std::max(n, 1)), sm(4 * (size_t)std::max(n, 1));
  int m = orc_range_filter(raw, stride, n, t_rel, rmax, rmin, sc.data());   // LiDARData ctor
  orc_transform_f32(sc.data(), m, T_LtoM, sm.data());                       // transform_to_map tail
  const int k = c->map_ngbr_num;
  std::vector<float> nb((size_t)std::max(m, 1) * k * 4), d2((size_t)std::max(m, 1) * k);
  std::vector<int> found(std::max(m, 1));
  ref_ikd_knn(tree, sm.data(), m, k, nb.data(), d2.data(), found.data(), threads); // get_neighbors -> Nearest_Search
  std::vector<uint8_t> ok(std::max(m, 1));
  for (int i = 0; i < m; i++)
    ok[i] = (found[i] >= k && !((double)d2[(size_t)i * k + k - 1] > c->map_ngbr_max_d)) ? 1 : 0; // LidarHelper.cpp:280-285
  return orc_scan_linsys_nb(c, P, sc.data(), m, nb.data(), ok.data(), HtH, Htr, cnt, Hc, rc, n_c, threads);
}
