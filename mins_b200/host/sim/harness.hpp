// Closed-loop IMU + LiDAR simulation in host C++ (BASELINE.json configs 0 and 4): the run_simulation loop of
// mins/src/run_simulation.cpp:71-130 restated - B-spline ground truth, IMU at 200 Hz, LiDAR at 10 Hz ray-cast against the floor
// plan ON THE DEVICE (ml_sim_raycast), clone window at 20 Hz / 1 s, and mins_gpu::UpdaterLidar (the C++ shim over the C ABI)
// doing every LiDAR update - plus the evaluator's ATE (mins_eval/src/functions/ResultTrajectory.cpp:70-97).
#pragma once
#include "../lidar_gpu.hpp"
#include "ekf.hpp"

#include <chrono>
#include <random>

namespace mins_sim {

// ---- Simulator::load_plane_data (mins/src/sim/Simulator.cpp:990-1096) -> rows of 22 doubles for ml_sim_set_world ---------------
// [A B C D | top_left | bottom_right | V1 V2 V3 V4] with the plane equation and edge directions of SimulationPlane.h:43-53,80-87
inline std::vector<double> load_floorplan_planes(const std::string &path) {
  std::ifstream f(path);
  if (!f)
    throw std::runtime_error("cannot open " + path);
  double units = 1.0;
  std::vector<std::vector<double>> segs;
  std::string line;
  while (std::getline(f, line)) {
    if (line.find("#") == 0) // only lines that START with the comment sign (`!current_line.find("#")`)
      continue;
    std::istringstream s(line);
    std::string field;
    std::vector<double> data;
    while (std::getline(s, field, ',')) {
      bool blank = true;
      for (char ch : field)
        if (!isspace((unsigned char)ch))
          blank = false;
      if (blank)
        continue;
      data.push_back(atof(field.c_str()));
    }
    if (data.size() == 1)
      units = data[0];
    if (data.size() == 4)
      segs.push_back(data);
  }
  if (segs.empty())
    throw std::runtime_error("no plane in " + path);
  std::vector<double> out;
  auto add = [&](const double *tl, const double *tr, const double *bl, const double *br) {
    double e1[3], e2[3], N[3];
    for (int i = 0; i < 3; i++)
      e1[i] = tr[i] - tl[i], e2[i] = bl[i] - tl[i];
    N[0] = e1[1] * e2[2] - e1[2] * e2[1], N[1] = e1[2] * e2[0] - e1[0] * e2[2], N[2] = e1[0] * e2[1] - e1[1] * e2[0];
    const double D = -(N[0] * tl[0] + N[1] * tl[1] + N[2] * tl[2]);
    out.insert(out.end(), {N[0], N[1], N[2], D, tl[0], tl[1], tl[2], br[0], br[1], br[2]});
    const double *from[4] = {tl, tl, br, br}, *to[4] = {tr, bl, tr, bl};
    for (int k = 0; k < 4; k++) {
      double v[3] = {to[k][0] - from[k][0], to[k][1] - from[k][1], to[k][2] - from[k][2]};
      const double n = norm3(v);
      out.insert(out.end(), {v[0] / n, v[1] / n, v[2] / n});
    }
  };
  double mnx = INFINITY, mxx = -INFINITY, mny = INFINITY, mxy = -INFINITY;
  for (auto &d : segs) {
    mnx = std::min({mnx, d[0], d[2]}), mxx = std::max({mxx, d[0], d[2]});
    mny = std::min({mny, d[1], d[3]}), mxy = std::max({mxy, d[1], d[3]});
    const double tl[3] = {units * d[0], units * d[1], 4}, tr[3] = {units * d[2], units * d[3], 4};
    const double bl[3] = {units * d[0], units * d[1], -1}, br[3] = {units * d[2], units * d[3], -1};
    add(tl, tr, bl, br);
  }
  for (double z : {-1.0, 4.0}) { // floor and ceiling (:1078-1090)
    const double tl[3] = {units * mnx, units * mny, z}, tr[3] = {units * mnx, units * mxy, z};
    const double bl[3] = {units * mxx, units * mny, z}, br[3] = {units * mxx, units * mxy, z};
    add(tl, tr, bl, br);
  }
  return out;
}

// ray bearings in the LiDAR frame, v-major like Simulator.cpp:756-757 (for v: for h), d = [sin(ph) cos(th), sin(ph) sin(th), cos(ph)],
// ph = 90 - v, th = h (:767-773)
inline std::vector<double> ray_dirs(const std::vector<double> &v_angles, const std::vector<double> &h_angles) {
  std::vector<double> d;
  d.reserve(3 * v_angles.size() * h_angles.size());
  for (double v : v_angles)
    for (double h : h_angles) {
      const double ph = M_PI / 180.0 * (90.0 - v), th = M_PI / 180.0 * h;
      d.insert(d.end(), {std::sin(ph) * std::cos(th), std::sin(ph) * std::sin(th), std::cos(ph)});
    }
  return d;
}

// ---- the StateView mins_gpu::UpdaterLidar needs, over the minimal EKF core --------------------------------------------------
struct SimStateView : mins_gpu::StateView {
  State *st;
  int calib_dummy = 0, dt_dummy = 0;
  explicit SimStateView(State *s) : st(s) {}
  bool clones_empty() const override { return st->clones.empty(); }
  size_t n_clones() const override { return st->clones.size(); }
  int intr_order() const override { return st->intr_order; }
  double time() const override { return st->time; }
  double oldest_clone_time() const override { return st->oldest_clone_time(); }
  double oldest_3rd_clone_time() const override { return st->oldest_3rd_clone_time(); }
  double newest_clone_time() const override { return st->newest_clone_time(); }
  double lidar_dt(int) const override { return 0.0; }
  const void *calib_var(int) const override { return &calib_dummy; }
  const void *dt_var(int) const override { return &dt_dummy; }
  void calibration(int, double RItoL[9], double pIinL[3]) const override {
    std::memcpy(RItoL, st->R_ItoL, sizeof(double) * 9);
    std::memcpy(pIinL, st->p_IinL, sizeof(double) * 3);
  }
  bool get_interpolated_jacobian(double t, int, double RGtoI[9], double pIinG[3], std::vector<mins_gpu::JacBlock> &dTdx) override {
    std::vector<State::JBlock> blocks;
    if (!st->get_interpolated_jacobian(t, RGtoI, pIinG, blocks))
      return false;
    dTdx.clear();
    for (auto &b : blocks) {
      mins_gpu::JacBlock jb;
      jb.var = &st->clones.at(b.t); // std::map nodes are stable: the clone's address identifies the state variable
      jb.width = 6;
      std::memcpy(jb.D, b.D, sizeof(jb.D));
      dTdx.push_back(jb);
    }
    return true;
  }
  bool get_interpolated_pose(double t, double RGtoI[9], double pIinG[3]) override { return st->get_interpolated_pose(t, RGtoI, pIinG); }
  bool have_clone(double t) const override { return st->have_clone(t); }
  double intr_ori_cov() const override { return st->intr_err_ori; }
  double intr_pos_cov() const override { return st->intr_err_pos; }
  std::vector<int> columns(const std::vector<const void *> &H_order, const std::vector<int> &widths) const {
    std::vector<int> idx;
    for (size_t k = 0; k < H_order.size(); k++) {
      const Clone *c = static_cast<const Clone *>(H_order[k]);
      for (int j = 0; j < widths[k]; j++)
        idx.push_back(c->idx + j);
    }
    return idx;
  }
  void get_marginal_covariance(const std::vector<const void *> &H_order, const std::vector<int> &widths, double *P) override {
    st->marginal_covariance(columns(H_order, widths), P);
  }
  bool last_update_ok = true;
  void EKFUpdate(const std::vector<const void *> &H_order, const std::vector<int> &widths, const double *H, const double *res, int rows,
                 int) override {
    last_update_ok = st->ekf_update(columns(H_order, widths), H, res, rows);
  }
};

struct SimConfig {
  unsigned long long seed = 0;
  double duration = 20.0;
  double freq_imu = 200.0, freq_lidar = 10.0, clone_freq = 20.0, window_size = 1.0; // config_simulation.yaml, config_estimator.yaml
  int intr_order = 3;
  double gyro_noise = 1.6968e-3, accel_noise = 2.0e-2, gyro_walk = 1.9393e-5, accel_walk = 3.0e-3; // config_imu.yaml
  double h_resolution = 2.0;
  double init_ori_err = 0.01, init_pos_err = 0.02;
  bool use_lidar = true, use_gt_lidar = true, no_noise = false;
  std::string bspline_path, floorplan_path, traj_out, state_log_out;
  mins_gpu::OptionsLidar op;
  int device = 0;
};

struct SimResult {
  double rmse_ori_deg = 0, rmse_pos_m = 0, nees_pos = 0, final_pos_err = 0;
  int n_updates = 0, n_scans = 0, n_clones = 0;
  long n_acc = 0, n_meas = 0;
  double wall_s = 0, ldr_wall_s = 0, raycast_wall_s = 0;
  int map_points = 0;
  std::vector<std::vector<double>> traj; // [t, p, q_jpl, 12 covariance entries]
};

struct PoseSample {
  double t, R[9], p[3];
  double cov[12] = {0}; // upper triangles of the orientation and the position block of the IMU pose covariance (State_Logger.h:333-339)
};

// ov_eval AlignTrajectory se3 + ResultTrajectory::calculate_ate (ResultTrajectory.cpp:70-97)
inline void ate(const std::vector<PoseSample> &est, const std::vector<PoseSample> &gt, double *rmse_ori_deg, double *rmse_pos) {
  const size_t n = est.size();
  double mu_e[3] = {0, 0, 0}, mu_g[3] = {0, 0, 0};
  for (size_t i = 0; i < n; i++)
    for (int a = 0; a < 3; a++)
      mu_e[a] += est[i].p[a] / n, mu_g[a] += gt[i].p[a] / n;
  double W[9] = {0};
  for (size_t i = 0; i < n; i++)
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++)
        W[3 * a + b] += (gt[i].p[a] - mu_g[a]) * (est[i].p[b] - mu_e[b]);
  double R[9], t[3], Rmu[3];
  kabsch(W, R);
  m3_vec(R, mu_e, Rmu);
  for (int a = 0; a < 3; a++)
    t[a] = mu_g[a] - Rmu[a];
  double so = 0, sp = 0;
  for (size_t i = 0; i < n; i++) {
    double pa[3], d[3], RIG[9], Rt[9], M[9], w[3];
    m3_vec(R, est[i].p, pa);
    for (int a = 0; a < 3; a++)
      d[a] = pa[a] + t[a] - gt[i].p[a];
    sp += d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
    m3_T(est[i].R, Rt);
    m3_mul(R, Rt, RIG); // R_est_ItoG aligned
    m3_mul(gt[i].R, RIG, M);
    log_so3(M, w);
    const double e = norm3(w) * 180.0 / M_PI;
    so += e * e;
  }
  *rmse_ori_deg = std::sqrt(so / n);
  *rmse_pos = std::sqrt(sp / n);
}

} // namespace mins_sim
