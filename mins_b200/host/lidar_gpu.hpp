// Host side of the B200 LiDAR update path, in the reference's own language (C++14, header only).
//
// Mirrors the operator interface of mins/src/update/lidar/{UpdaterLidar,LidarHelper,LidarTypes}.h — same class and
// method names, same argument meaning, same bool/exit error behaviour — with every numeric step delegated to the C ABI
// of include/mins_lidar.h. The reference's types that cannot be compiled in this image (Eigen matrices, PCL clouds,
// mins::State) are replaced by plain arrays and by the abstract StateView below; INTEGRATION.md shows the ~40-line
// StateView a MINS maintainer writes over mins::State / StateHelper, after which SystemManager.cpp:70,105,182-183 use
// this UpdaterLidar unchanged.
//
// The queueing policy (three stacks raw -> new -> used, clone-window tests, +0.05 s delay, 1 s staleness re-init,
// UpdaterLidar.cpp:50-145) stays on the host exactly as in the reference.
//
// The C ABI is reached through a table of function pointers (ml_api) so that the policy can be unit-tested on a CPU-only
// box against a recording mock (tests/test_host_shim.py); ml_api::real() binds the symbols of libmins_lidar.so.
#pragma once
#include "../../include/mins_lidar.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <algorithm>
#include <memory>
#include <vector>

namespace mins_gpu {

// ---- the C ABI as a table --------------------------------------------------------------------------------------
struct ml_api {
  decltype(&ml_create) create;
  decltype(&ml_destroy) destroy;
  decltype(&ml_last_error) last_error;
  decltype(&ml_map_reset) map_reset;
  decltype(&ml_map_initialized) map_initialized;
  decltype(&ml_map_build_from_scan) map_build_from_scan;
  decltype(&ml_map_add_scan) map_add_scan;
  decltype(&ml_map_reanchor) map_reanchor;
  decltype(&ml_map_flatten) map_flatten;
  decltype(&ml_map_size) map_size;
  decltype(&ml_scan_upload) scan_upload;
  decltype(&ml_scan_deskew) scan_deskew;
  decltype(&ml_scan_downsample) scan_downsample;
  decltype(&ml_scan_transform) scan_transform;
  decltype(&ml_scan_update) scan_update;
  decltype(&ml_info_to_sqrt) info_to_sqrt;
  decltype(&ml_linsys) linsys;
  decltype(&ml_icp_default_options) icp_default_options;
  decltype(&ml_scan_icp) scan_icp;
  decltype(&ml_host_alloc) host_alloc;
  decltype(&ml_host_free) host_free;
  decltype(&ml_map_build) map_build;
  decltype(&ml_map_add) map_add;
#ifndef MINS_GPU_NO_LINK
  static ml_api real() {
    return ml_api{&ml_create,      &ml_destroy,         &ml_last_error,  &ml_map_reset,   &ml_map_initialized, &ml_map_build_from_scan,
                  &ml_map_add_scan, &ml_map_reanchor,   &ml_map_flatten, &ml_map_size,    &ml_scan_upload,     &ml_scan_deskew,
                  &ml_scan_downsample, &ml_scan_transform, &ml_scan_update, &ml_info_to_sqrt, &ml_linsys, &ml_icp_default_options,
                  &ml_scan_icp, &ml_host_alloc, &ml_host_free, &ml_map_build, &ml_map_add};
  }
#endif
};

// ---- OptionsLidar (mins/src/options/OptionsLidar.h:38-99), the fields the path reads ----------------------------
struct OptionsLidar {
  int max_n = 2;
  double max_range = 100.0, min_range = 0.05, chi2_mult = 1.0;
  bool raw_do_downsample = true, raw_remove_motion_blur = false;
  double raw_downsample_size = 0.1, raw_noise = 0.01, raw_point_dt = 1e-6;
  int raw_total_points = 0; // v_angles.size() * h_angles.size() of the sensor (LidarHelper.cpp:57): rays of one sweep, for the
                            // de-skew time stamps; 0 = the size of the scan
  int deskew_knots = 16;    // poses sampled over the sweep for ml_scan_deskew (the device blends the two bracketing knots)
  bool map_do_downsample = true, map_use_icp = false;
  double map_downsample_size = 0.1, map_noise = 0.1, map_ngbr_max_d = 0.5, map_decay_time = 30, map_decay_dist = 100;
  double map_icp_dist = 20; // OptionsLidar.h:97
  unsigned long long icp_seed = 0; // the reference seeds its ICP samplers from std::random_device; here it is explicit
  int map_ngbr_num = 5;
  double plane_max_p2pd = 0.1, plane_max_condi = 100.0;
  bool do_calib_ext = false, do_calib_dt = false;
};

// d(interpolated pose)/d(state variable): State::get_interpolated_jacobian's dTdx[j] with order[j] (State.cpp:745-816).
// `var` identifies the state variable (the reference keys map_hx by shared_ptr<ov_type::Type>; any stable id works).
struct JacBlock {
  const void *var;
  int width; // 6 = pose, 1 = time offset
  double D[36];
};

// What the path needs from the untouched EKF core. One implementation forwards to mins::State / StateHelper.
struct StateView {
  virtual ~StateView() {}
  virtual bool clones_empty() const = 0;                 // state->clones.empty()
  virtual size_t n_clones() const = 0;                   // state->clones.size()
  virtual int intr_order() const = 0;                    // state->op->intr_order
  virtual double time() const = 0;                       // state->time
  virtual double oldest_clone_time() const = 0;          // State.cpp:950
  virtual double oldest_3rd_clone_time() const = 0;
  virtual double newest_clone_time() const = 0;          // State.cpp:986
  virtual double lidar_dt(int id) const = 0;             // state->lidar_dt.at(id)->value()(0)
  virtual const void *calib_var(int id) const = 0;       // state->lidar_extrinsic.at(id)
  virtual const void *dt_var(int id) const = 0;          // state->lidar_dt.at(id)
  virtual void calibration(int id, double RItoL[9], double pIinL[3]) const = 0;
  // State::get_interpolated_jacobian(t, R, p, "LIDAR", id, dTdx, order) — fej values (State.cpp:675-826)
  virtual bool get_interpolated_jacobian(double t, int id, double RGtoI[9], double pIinG[3], std::vector<JacBlock> &dTdx) = 0;
  // State::get_interpolated_pose(t, R, p) — current estimate (State.cpp:597-673)
  virtual bool get_interpolated_pose(double t, double RGtoI[9], double pIinG[3]) = 0;
  virtual bool have_clone(double t) const = 0;           // State.cpp:1243
  virtual double intr_ori_cov() const = 0;               // state->intr_ori_cov(clone_freq, intr_order), State.cpp:1238
  virtual double intr_pos_cov() const = 0;
  // StateHelper::get_marginal_covariance(state, H_order) (StateHelper.cpp:227-254): row-major n x n
  virtual void get_marginal_covariance(const std::vector<const void *> &H_order, const std::vector<int> &widths, double *P) = 0;
  // StateHelper::EKFUpdate(state, H_order, H, res, R = I) (StateHelper.cpp:121-200): H is rows x total_hx row-major
  virtual void EKFUpdate(const std::vector<const void *> &H_order, const std::vector<int> &widths, const double *H, const double *res,
                         int rows, int total_hx) = 0;
};

// ---- small fixed-size algebra ------------------------------------------------------------------------------------
inline void mul33(const double *A, const double *B, double *C) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
inline void tr33(const double *A, double *T) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      T[3 * i + j] = A[3 * j + i];
}
inline void mv3(const double *A, const double *v, double *o) {
  for (int i = 0; i < 3; i++)
    o[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
}

// ---- LiDARData (LidarTypes.h:37-66): a queued scan. The clouds live on the device once the scan is being processed;
// while queued only the raw host copy is kept (the range filter of the constructor runs on the device at upload). ----
// The raw points sit in a pinned, device-mapped buffer from the updater's pool, PointXYZ layout (x, y, z, pad: stride 4, the
// reference's own 16-byte points), so that the device reads them in place when the scan's turn comes.
struct LiDARData {
  double time = -1;
  int id = -1;
  float *xyz = nullptr; // n x 4, sensor frame (owned by the updater's pool)
  int n = 0;
  int size() const { return n; }
};

// ---- KD_TREE (ikd_Tree.h:92-160) as MINS code OUTSIDE the updater sees it: Initializer.cpp:191-201,330-338 asks
// `tree->initialized()`, ROSPublisher.cpp:432-470 calls `tree->flatten(tree->Root_Node, pts, NOT_RECORD)`. A facade over one
// ml_handle with the same member names, so those call sites compile unchanged; the points are PointXYZI's four fields.
struct PointXYZI {
  float x, y, z, intensity;
};
typedef PointXYZI PointType;
typedef std::vector<PointType> PointVector;
enum delete_point_storage_set { NOT_RECORD, DELETE_POINTS_REC, MULTI_THREAD_REC };
struct KD_TREE_NODE; // opaque: Root_Node is only ever handed back to flatten()
struct KD_TREE {
  ml_api api;
  ml_handle *h = nullptr;
  float downsample_size = 0.3f;
  KD_TREE_NODE *Root_Node = nullptr; // non-null while the map is initialised (refreshed by every member below)
  KD_TREE(const ml_api &a, ml_handle *handle) : api(a), h(handle) {}
  bool initialized() { return refresh(); }                                  // ikd_Tree.h `initialized()`
  void reset() {                                                            // KD_TREE::reset
    api.map_reset(h);
    refresh();
  }
  void set_downsample_param(float ds) { downsample_size = ds; }             // ikd_Tree.cpp:125-127
  void Build(const PointVector &pts) {                                      // KD_TREE::Build (ikd_Tree.cpp:344-356)
    api.map_build(h, pts.empty() ? nullptr : &pts[0].x, (int)pts.size());
    refresh();
  }
  int Add_Points(PointVector &pts, bool downsample_on) {                    // KD_TREE::Add_Points (ikd_Tree.cpp:401-479)
    int n_live = 0;
    api.map_add(h, pts.empty() ? nullptr : &pts[0].x, (int)pts.size(), downsample_on ? 1 : 0, downsample_size, &n_live);
    refresh();
    return n_live;
  }
  void flatten(KD_TREE_NODE *, PointVector &Storage, delete_point_storage_set) { // KD_TREE::flatten (ikd_Tree.cpp:1359-1386)
    int n = 0;
    api.map_size(h, &n);
    Storage.resize((size_t)std::max(n, 0));
    if (n > 0)
      api.map_flatten(h, &Storage[0].x, n, &n);
  }
  int size() {
    int n = 0;
    api.map_size(h, &n);
    return n;
  }
  int validnum() { return size(); }

private:
  bool refresh() {
    const bool init = h && api.map_initialized(h) > 0;
    Root_Node = init ? reinterpret_cast<KD_TREE_NODE *>(h) : nullptr;
    return init;
  }
};

// ---- iKDDATA (LidarTypes.h:71-81): one local map = one ml_handle ---------------------------------------------------
struct iKDDATA {
  int id = -1;
  double time = -1;         // anchor time of the map
  double last_up_time = -1; // last successful registration
  std::shared_ptr<KD_TREE> tree; // tree->h is the handle the updater drives
};

// statistics the reference keeps in UpdaterStatistics (UpdaterStatistics.cpp:47-83), accumulated from ml_counters
struct UpdaterStatistics {
  long n_meas = 0, n_acp = 0, n_rej = 0;
  double sum_chi = 0, sum_chi2 = 0, sum_res2 = 0;
  void add(const ml_counters &c) {
    n_acp += c.n_acc;
    n_rej += c.n_chi_rej;
    n_meas += c.n_acc + c.n_chi_rej;
    sum_chi += c.sum_chi, sum_chi2 += c.sum_chi2, sum_res2 += c.sum_res2;
  }
};

// ---- LidarHelper (LidarHelper.h:45-70): the pieces that are host arithmetic; the rest is one ABI call each --------
struct LidarHelper {
  // LidarHelper::transform_to_map's pose algebra (LidarHelper.cpp:141-151): T_LtoM as fp32 row-major 4x4
  static void T_LtoM(const double RGtoL[9], const double pLinG[3], const double RGtoM[9], const double pMinG[3], float T[16]) {
    double RLtoG[9], RLtoM[9], d[3], p[3];
    tr33(RGtoL, RLtoG);
    mul33(RGtoM, RLtoG, RLtoM);
    for (int i = 0; i < 3; i++)
      d[i] = pLinG[i] - pMinG[i];
    mv3(RGtoM, d, p);
    for (int i = 0; i < 16; i++)
      T[i] = (i % 5 == 0) ? 1.f : 0.f;
    for (int i = 0; i < 3; i++) {
      for (int j = 0; j < 3; j++)
        T[4 * i + j] = (float)RLtoM[3 * i + j];
      T[4 * i + 3] = (float)p[i];
    }
  }
  // pose of the LiDAR at time t: R_GtoL = R_ItoL R_GtoI, p_LinG = p_IinG + R_GtoI^T p_LinI (UpdaterLidar::get_pose_LinG)
  static bool get_pose_LinG(StateView &st, int id, double t, double RGtoL[9], double pLinG[3]) {
    double RGtoI[9], pIinG[3], RItoL[9], pIinL[3], RLtoI[9], pLinI[3], RItoG[9], tmp[3];
    if (!st.get_interpolated_pose(t, RGtoI, pIinG))
      return false;
    st.calibration(id, RItoL, pIinL);
    tr33(RItoL, RLtoI);
    mv3(RLtoI, pIinL, pLinI);
    for (int i = 0; i < 3; i++)
      pLinI[i] = -pLinI[i];
    mul33(RItoL, RGtoI, RGtoL);
    tr33(RGtoI, RItoG);
    mv3(RItoG, pLinI, tmp);
    for (int i = 0; i < 3; i++)
      pLinG[i] = pIinG[i] + tmp[i];
    return true;
  }
};

} // namespace mins_gpu

#include "updater_lidar_gpu.hpp"
