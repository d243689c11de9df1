// CPU-only test driver for the host shim's queueing policy (UpdaterLidar.cpp:50-160) against a recording mock of the C
// ABI and a scripted StateView. Prints one line per ABI call / decision; tests/test_host_shim.py checks the trace.
// Built with -DMINS_GPU_NO_LINK (no CUDA library needed).
#define MINS_GPU_NO_LINK
#include "lidar_gpu.hpp"

#include <cstdio>
#include <cstring>
#include <string>

using namespace mins_gpu;

struct ml_handle {
  int id;
  bool init = false;
  int n_map = 0;
};
static int g_next = 0;
static int g_update_rc = ML_OK;

static int m_create(const ml_config *, ml_handle **out) {
  *out = new ml_handle();
  (*out)->id = g_next++;
  std::printf("create h%d\n", (*out)->id);
  return ML_OK;
}
static void m_destroy(ml_handle *h) { delete h; }
static const char *m_err(const ml_handle *) { return "mock"; }
static int m_reset(ml_handle *h) {
  h->init = false;
  return ML_OK;
}
static int m_inited(const ml_handle *h) { return h->init ? 1 : 0; }
static int m_build(ml_handle *h) {
  h->init = true;
  std::printf("h%d map_build_from_scan\n", h->id);
  return ML_OK;
}
static int m_add_scan(ml_handle *h, int ds, float size, int *) {
  std::printf("h%d map_add_scan ds=%d size=%.2f\n", h->id, ds, size);
  return ML_OK;
}
static int m_reanchor(ml_handle *h, const double T[16], double now, double dtime, double ddist, int ds, float size, int *) {
  std::printf("h%d map_reanchor tx=%.3f now=%.3f decay=%.1f/%.1f ds=%d size=%.2f\n", h->id, T[3], now, dtime, ddist, ds, size);
  return ML_OK;
}
static int m_flatten(ml_handle *h, float *out, int cap, int *n) {
  for (int i = 0; i < cap && i < h->n_map; i++)
    out[4 * i] = (float)i, out[4 * i + 1] = out[4 * i + 2] = 0.f, out[4 * i + 3] = 0.25f;
  if (n)
    *n = h->n_map;
  return ML_OK;
}
static int m_size(ml_handle *h, int *n) {
  *n = h->n_map;
  return ML_OK;
}
static int m_map_build(ml_handle *h, const float *, int n) {
  h->init = n > 0;
  h->n_map = n;
  std::printf("h%d map_build n=%d\n", h->id, n);
  return ML_OK;
}
static int m_map_add(ml_handle *h, const float *, int n, int ds, float size, int *n_live) {
  h->n_map += n;
  if (n_live)
    *n_live = h->n_map;
  std::printf("h%d map_add n=%d downsample=%d size=%.2f\n", h->id, n, ds, size);
  return ML_OK;
}
static int m_upload(ml_handle *h, const float *, int stride, int n, float t_rel, float rmin, float rmax, int *) {
  std::printf("h%d scan_upload n=%d stride=%d t_rel=%.3f range=[%.2f,%.1f]\n", h->id, n, stride, t_rel, rmin, rmax);
  return ML_OK;
}
static int m_deskew(ml_handle *h, const float *knots, int n_knots, int total) {
  std::printf("h%d scan_deskew knots=%d total=%d tx_first=%.4f tx_last=%.4f\n", h->id, n_knots, total, knots[3], knots[(n_knots - 1) * 12 + 3]);
  return ML_OK;
}
static int g_icp_fail = 0; // the next g_icp_fail ICP calls report failure (max iterations reached)
static int m_down(ml_handle *h, float leaf, int *) {
  std::printf("h%d scan_downsample leaf=%.2f\n", h->id, leaf);
  return ML_OK;
}
static int m_transform(ml_handle *h, const float T[16]) {
  std::printf("h%d scan_transform tx=%.3f\n", h->id, T[3]);
  return ML_OK;
}
static int m_update(ml_handle *h, const float *, int stride, int n, float t_rel, float, float, float leaf, const float T[16],
                    const ml_scan_consts *c, const double *P, double *HtH, double *Htr, ml_counters *cnt) {
  std::printf("h%d scan_update n=%d stride=%d t_rel=%.3f leaf=%.2f tx=%.3f hx=%d nI=%d nA=%d calib_col=%d cols=[", h->id, n, stride, t_rel,
              leaf, T[3], c->total_hx, c->nI, c->nA, c->calib_col);
  for (int j = 0; j < c->nI; j++)
    std::printf("%d ", c->dTdxI[j].col);
  std::printf("|");
  for (int j = 0; j < c->nA; j++)
    std::printf(" %d", c->dTdxA[j].col);
  std::printf("] at_clone=%d%d k=%d chi2=%.4f P00=%.3g\n", c->at_clone_m, c->at_clone_a, c->map_ngbr_num, c->chi2_mult * 7.814727903251179, P[0]);
  for (int i = 0; i < c->total_hx * c->total_hx; i++)
    HtH[i] = 0;
  for (int i = 0; i < c->total_hx; i++) {
    HtH[i * c->total_hx + i] = 4.0;
    Htr[i] = 2.0;
  }
  cnt->n_acc = g_update_rc == ML_OK ? n : 0;
  cnt->n_chi_rej = 1;
  return g_update_rc;
}
static int m_linsys(ml_handle *h, const ml_scan_consts *c, const double *P, double *HtH, double *Htr, ml_counters *cnt) {
  std::printf("h%d linsys hx=%d\n", h->id, c->total_hx);
  for (int i = 0; i < c->total_hx * c->total_hx; i++)
    HtH[i] = 0;
  for (int i = 0; i < c->total_hx; i++) {
    HtH[i * c->total_hx + i] = 4.0;
    Htr[i] = 2.0;
  }
  cnt->n_acc = 10;
  return ML_OK;
}
static void m_icp_defaults(ml_icp_options *o) {
  std::memset(o, 0, sizeof(*o));
  o->max_iter = 40;
}
static int m_icp(ml_handle *h, const float T[16], const ml_icp_options *o, ml_icp_result *r) {
  std::printf("h%d scan_icp tx=%.3f max_dist=%.1f seed=%llu\n", h->id, T[3], o->max_dist, (unsigned long long)o->seed);
  std::memset(r, 0, sizeof(*r));
  for (int i = 0; i < 16; i++)
    r->T_LtoM[i] = r->T_icp[i] = T[i];
  if (g_icp_fail > 0) { // lidar->icp_success = false: T_LtoM stays the initial guess (LidarHelper.cpp:177-179)
    g_icp_fail--;
    r->success = 0, r->max_iter_reached = 1;
    return ML_OK;
  }
  r->T_LtoM[3] += 0.5f; // a visible correction
  r->success = 1;
  return ML_OK;
}
static int g_host_allocs = 0;
static void *m_host_alloc(size_t bytes) {
  g_host_allocs++;
  return std::malloc(bytes);
}
static void m_host_free(void *p) { std::free(p); }
static int m_sqrt(const double *HtH, const double *Htr, int n, double *Rc, double *rc) {
  for (int i = 0; i < n * n; i++)
    Rc[i] = 0;
  for (int i = 0; i < n; i++) {
    Rc[i * n + i] = std::sqrt(HtH[i * n + i]);
    rc[i] = Htr[i] / Rc[i * n + i];
  }
  return n;
}

// clones every 0.05 s; cubic interpolation touches 4 clones (intr_order 3)
struct MockState : StateView {
  std::vector<double> clones;
  double now = 0, dt = 0.0;
  int ids[64];
  int calib_id, dt_id;
  bool clones_empty() const override { return clones.empty(); }
  size_t n_clones() const override { return clones.size(); }
  int intr_order() const override { return 3; }
  double time() const override { return now; }
  double oldest_clone_time() const override { return clones.empty() ? -INFINITY : clones.front(); } // State.cpp:950-959
  double oldest_3rd_clone_time() const override { return clones.size() < 3 ? -INFINITY : clones[2]; }
  double newest_clone_time() const override { return clones.empty() ? -INFINITY : clones.back(); }
  double lidar_dt(int) const override { return dt; }
  const void *calib_var(int) const override { return &calib_id; }
  const void *dt_var(int) const override { return &dt_id; }
  void calibration(int, double R[9], double p[3]) const override {
    for (int i = 0; i < 9; i++)
      R[i] = (i % 4 == 0);
    p[0] = p[1] = p[2] = 0;
  }
  int bracket(double t) const { // index of the clone at or before t
    int k = -1;
    for (size_t i = 0; i < clones.size(); i++)
      if (clones[i] <= t + 1e-9)
        k = (int)i;
    return k;
  }
  bool get_interpolated_jacobian(double t, int, double R[9], double p[3], std::vector<JacBlock> &J) override {
    int k = bracket(t);
    J.clear();
    if (have_clone(t)) { // exactly at a clone: the pose is that clone, one identity block (State.cpp:690-700)
      get_interpolated_pose(t, R, p);
      JacBlock b;
      b.var = &ids[k];
      b.width = 6;
      for (int q = 0; q < 36; q++)
        b.D[q] = (q % 7 == 0) ? 1.0 : 0.0;
      J.push_back(b);
      return true;
    }
    if (k < 1 || k + 2 >= (int)clones.size())
      return false;
    get_interpolated_pose(t, R, p);
    for (int j = k - 1; j <= k + 2; j++) {
      JacBlock b;
      b.var = &ids[j];
      b.width = 6;
      for (int q = 0; q < 36; q++)
        b.D[q] = (q % 7 == 0) ? 0.25 : 0.0;
      J.push_back(b);
    }
    return true;
  }
  bool get_interpolated_pose(double t, double R[9], double p[3]) override {
    if (clones.empty() || t < clones.front() - 1e-9 || t > clones.back() + 1e-9)
      return false;
    for (int i = 0; i < 9; i++)
      R[i] = (i % 4 == 0);
    p[0] = t, p[1] = 0, p[2] = 0; // 1 m/s along x
    return true;
  }
  bool have_clone(double t) const override {
    for (double c : clones)
      if (std::fabs(c - t) < 1e-9)
        return true;
    return false;
  }
  double intr_ori_cov() const override { return 1e-6; }
  double intr_pos_cov() const override { return 1e-5; }
  void get_marginal_covariance(const std::vector<const void *> &, const std::vector<int> &w, double *P) override {
    int n = 0;
    for (int x : w)
      n += x;
    for (int i = 0; i < n * n; i++)
      P[i] = (i % (n + 1) == 0) ? 1e-4 : 0.0;
  }
  void EKFUpdate(const std::vector<const void *> &order, const std::vector<int> &, const double *H, const double *res, int rows,
                 int hx) override {
    std::printf("EKFUpdate vars=%zu rows=%d hx=%d H00=%.2f r0=%.2f\n", order.size(), rows, hx, H[0], res[0]);
  }
};

int main() {
  ml_api api{m_create, m_destroy, m_err,    m_reset, m_inited, m_build,     m_add_scan, m_reanchor,
             m_flatten, m_size,   m_upload, m_deskew, m_down,  m_transform, m_update,   m_sqrt,
             m_linsys,  m_icp_defaults, m_icp, m_host_alloc, m_host_free, m_map_build, m_map_add};
  OptionsLidar op;
  op.max_n = 1, op.raw_do_downsample = true, op.raw_downsample_size = 0.3, op.map_downsample_size = 0.3, op.min_range = 1.0;
  op.do_calib_ext = true;
  MockState st;
  UpdaterLidar up(&st, op, api, 1024, 4096);
  std::vector<float> pts(3 * 10, 2.0f);
  auto state_of = [&]() { std::printf("stacks raw=%zu new=%zu used=%zu\n", up.stack_lidar_raw.size(), up.stack_lidar_new.size(), up.stack_lidar_used.size()); };

  std::printf("# no clones yet: the scan waits in the raw stack\n");
  up.feed_measurement(0, 10.00, pts.data(), 10);
  state_of();
  std::printf("# clones 10.00..10.10: scan at 10.00 is old enough (10.00+0.05 <= 10.10): map is initialised from it\n");
  st.clones = {10.00, 10.05, 10.10};
  st.now = 10.10;
  up.feed_measurement(0, 10.10, pts.data(), 10);
  state_of();
  std::printf("# scan at 10.10 still newer than newest clone - 0.05: waits; more clones arrive\n");
  st.clones = {10.00, 10.05, 10.10, 10.15, 10.20, 10.25, 10.30};
  st.now = 10.30;
  up.feed_measurement(0, 10.12, pts.data(), 10);
  state_of();
  std::printf("# try_update: 10.10 is exactly a clone, 10.12 interpolates; both update then both register, in order\n");
  up.try_update();
  state_of();
  std::printf("chi n_acp=%ld n_rej=%ld last_up=%.2f anchor=%.2f\n", up.Chi[0].n_acp, up.Chi[0].n_rej, up.ikd_data[0]->last_up_time,
              up.ikd_data[0]->time);
  std::printf("# map anchor (10.00) is not newer than the 3rd-oldest clone (10.10): re-anchor to the newest clone\n");
  up.propagate_map_frame();
  std::printf("anchor=%.2f last_up=%.2f\n", up.ikd_data[0]->time, up.ikd_data[0]->last_up_time);
  up.propagate_map_frame(); // now the anchor is the newest clone: nothing to do
  std::printf("# a scan the device rejects entirely (no rows) stays in the new stack\n");
  up.feed_measurement(0, 10.20, pts.data(), 10);
  g_update_rc = ML_SOFT_NO_ROWS;
  up.try_update();
  state_of();
  g_update_rc = ML_OK;
  std::printf("# window slides past it: dropped by erase_invalid_dataset\n");
  st.clones = {10.25, 10.30, 10.35, 10.40, 10.45, 10.50};
  st.now = 10.50;
  up.feed_measurement(0, 10.52, pts.data(), 10);
  state_of();
  std::printf("# no registration for more than 1 s: the next usable scan re-initialises the map\n");
  st.clones = {11.40, 11.45, 11.50, 11.55, 11.60, 11.65};
  st.now = 11.65;
  up.feed_measurement(0, 11.50, pts.data(), 10);
  state_of();
  std::printf("anchor=%.2f last_up=%.2f t_hist=%zu\n", up.ikd_data[0]->time, up.ikd_data[0]->last_up_time, up.t_hist[0].size());
  std::printf("# map_use_icp: upload -> downsample -> ICP -> transform with the ICP result -> linsys; registration uses it too\n");
  up.op.map_use_icp = true, up.op.map_icp_dist = 50.0;
  up.feed_measurement(0, 11.56, pts.data(), 10);
  up.try_update();
  state_of();
  std::printf("icp_success=%d\n", (int)up.last_icp_success);
  std::printf("# ICP fails in update(): register_scan retries transform_to_map (LidarHelper.cpp:198); the retry succeeds -> registered with the retry's transform\n");
  st.clones = {11.50, 11.55, 11.60, 11.65, 11.70, 11.75, 11.80};
  st.now = 11.80;
  g_icp_fail = 1;
  up.feed_measurement(0, 11.62, pts.data(), 10);
  up.try_update();
  std::printf("icp_success=%d last_up=%.2f\n", (int)up.last_icp_success, up.ikd_data[0]->last_up_time);
  std::printf("# ICP fails twice: the scan is NOT added and last_up_time is NOT advanced (LidarHelper.cpp:201-206)\n");
  g_icp_fail = 2;
  up.feed_measurement(0, 11.66, pts.data(), 10);
  up.try_update();
  state_of();
  std::printf("icp_success=%d last_up=%.2f\n", (int)up.last_icp_success, up.ikd_data[0]->last_up_time);
  std::printf("# raw_remove_motion_blur: knots over the sweep from the state, ml_scan_deskew between upload and downsample\n");
  up.op.map_use_icp = false, up.op.raw_remove_motion_blur = true, up.op.raw_total_points = 1000, up.op.raw_point_dt = 1e-5, up.op.deskew_knots = 5;
  up.feed_measurement(0, 11.70, pts.data(), 10);
  up.try_update();
  state_of();
  std::printf("host_allocs=%d\n", g_host_allocs);
  std::printf("# KD_TREE facade: the members Initializer.cpp:191-201,330-338 and ROSPublisher.cpp:432-470 use on ikd->tree\n");
  {
    auto tree = up.ikd_data[0]->tree;
    tree->reset();
    std::printf("initialized=%d root=%d\n", (int)tree->initialized(), tree->Root_Node != nullptr);
    PointVector cloud(7, PointXYZI{1.f, 2.f, 3.f, 0.5f});
    tree->set_downsample_param(0.25f);
    tree->Build(cloud);
    std::printf("initialized=%d root=%d size=%d validnum=%d\n", (int)tree->initialized(), tree->Root_Node != nullptr, tree->size(), tree->validnum());
    PointVector more(3, PointXYZI{4.f, 5.f, 6.f, 0.75f});
    const int live = tree->Add_Points(more, true);
    PointVector pts_out;
    tree->flatten(tree->Root_Node, pts_out, NOT_RECORD);
    std::printf("live=%d flattened=%zu x_last=%.1f intensity=%.2f\n", live, pts_out.size(), pts_out.back().x, pts_out.back().intensity);
  }
  return 0;
}
