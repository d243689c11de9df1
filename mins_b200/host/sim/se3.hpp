// SO(3)/SE(3) helpers and the ground-truth B-spline of the closed-loop harness (host C++14, fp64).
// Conventions follow ov_core's quat_ops.h (thirdparty/open_vins/ov_core/src/utils/quat_ops.h): JPL quaternions,
// R = R_GtoI, error state R_true = exp_so3(-dth) R_hat. The spline restates thirdparty/open_vins/ov_core/src/sim/BsplineSE3.cpp
// (feed_trajectory :24-100, get_pose :102-136), which the MINS simulator uses for its ground truth.
#pragma once
#include "mat.hpp"

#include <fstream>
#include <map>
#include <sstream>
#include <string>

namespace mins_sim {

// quat_ops.h exp_so3 (Rodrigues)
inline void exp_so3(const double *w, double *R) {
  const double th = norm3(w);
  double K[9], KK[9];
  skew(w, K);
  m3_mul(K, K, KK);
  double a, b;
  if (th < 1e-7) {
    a = 1.0, b = 0.5;
  } else {
    a = std::sin(th) / th, b = (1.0 - std::cos(th)) / (th * th);
  }
  for (int i = 0; i < 9; i++)
    R[i] = ((i % 4 == 0) ? 1.0 : 0.0) + a * K[i] + b * KK[i];
}
// quat_ops.h log_so3: rotation vector of R
inline void log_so3(const double *R, double *w) {
  double c = 0.5 * (R[0] + R[4] + R[8] - 1.0);
  c = std::min(1.0, std::max(-1.0, c));
  const double th = std::acos(c);
  const double v[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
  if (th < 1e-7) {
    for (int i = 0; i < 3; i++)
      w[i] = 0.5 * v[i];
    return;
  }
  if (M_PI - th < 1e-6) { // near pi: use the symmetric part
    double A[9];
    for (int i = 0; i < 9; i++)
      A[i] = 0.5 * (R[i] + ((i % 4 == 0) ? 1.0 : 0.0));
    double ax[3] = {std::sqrt(std::max(A[0], 0.0)), std::sqrt(std::max(A[4], 0.0)), std::sqrt(std::max(A[8], 0.0))};
    int k = 0;
    for (int i = 1; i < 3; i++)
      if (ax[i] > ax[k])
        k = i;
    double col[3] = {A[k] / ax[k], A[3 + k] / ax[k], A[6 + k] / ax[k]};
    if (col[0] * v[0] + col[1] * v[1] + col[2] * v[2] < 0)
      for (int i = 0; i < 3; i++)
        col[i] = -col[i];
    const double n = norm3(col);
    for (int i = 0; i < 3; i++)
      w[i] = th * col[i] / n;
    return;
  }
  const double f = th / (2.0 * std::sin(th));
  for (int i = 0; i < 3; i++)
    w[i] = f * v[i];
}
// left Jacobian of SO(3) (quat_ops.h Jl_so3); Jr(w) = Jl(-w)
inline void Jl_so3(const double *w, double *J) {
  const double th = norm3(w);
  double K[9], KK[9];
  skew(w, K);
  m3_mul(K, K, KK);
  double a, b;
  if (th < 1e-7) {
    a = 0.5, b = 1.0 / 6.0;
  } else {
    a = (1.0 - std::cos(th)) / (th * th), b = (th - std::sin(th)) / (th * th * th);
  }
  for (int i = 0; i < 9; i++)
    J[i] = ((i % 4 == 0) ? 1.0 : 0.0) + a * K[i] + b * KK[i];
}
// quat_ops.h quat_2_Rot: JPL quaternion [x, y, z, w] -> rotation matrix (R_GtoI for q_GtoI)
inline void quat_jpl_to_rot(const double *qin, double *R) {
  double q[4];
  const double n = std::sqrt(qin[0] * qin[0] + qin[1] * qin[1] + qin[2] * qin[2] + qin[3] * qin[3]);
  for (int i = 0; i < 4; i++)
    q[i] = qin[i] / n;
  double S[9];
  skew(q, S);
  const double w = q[3];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      R[3 * i + j] = ((i == j) ? (2 * w * w - 1) : 0.0) - 2 * w * S[3 * i + j] + 2 * q[i] * q[j];
}
// quat_ops.h rot_2_quat
inline void rot_to_quat_jpl(const double *R, double *q) {
  const double T = R[0] + R[4] + R[8];
  if (R[0] >= T && R[0] >= R[4] && R[0] >= R[8]) {
    q[0] = std::sqrt((1 + 2 * R[0] - T) / 4);
    q[1] = (R[1] + R[3]) / (4 * q[0]), q[2] = (R[2] + R[6]) / (4 * q[0]), q[3] = (R[5] - R[7]) / (4 * q[0]);
  } else if (R[4] >= T && R[4] >= R[0] && R[4] >= R[8]) {
    q[1] = std::sqrt((1 + 2 * R[4] - T) / 4);
    q[0] = (R[1] + R[3]) / (4 * q[1]), q[2] = (R[5] + R[7]) / (4 * q[1]), q[3] = (R[6] - R[2]) / (4 * q[1]);
  } else if (R[8] >= T && R[8] >= R[0] && R[8] >= R[4]) {
    q[2] = std::sqrt((1 + 2 * R[8] - T) / 4);
    q[0] = (R[2] + R[6]) / (4 * q[2]), q[1] = (R[5] + R[7]) / (4 * q[2]), q[3] = (R[1] - R[3]) / (4 * q[2]);
  } else {
    q[3] = std::sqrt((1 + T) / 4);
    q[0] = (R[5] - R[7]) / (4 * q[3]), q[1] = (R[6] - R[2]) / (4 * q[3]), q[2] = (R[1] - R[3]) / (4 * q[3]);
  }
  if (q[3] < 0)
    for (int i = 0; i < 4; i++)
      q[i] = -q[i];
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (int i = 0; i < 4; i++)
    q[i] /= n;
}

// 4x4 rigid transforms as {R[9], t[3]}
struct SE3 {
  double R[9], t[3];
};
inline SE3 se3_mul(const SE3 &A, const SE3 &B) {
  SE3 C;
  m3_mul(A.R, B.R, C.R);
  m3_vec(A.R, B.t, C.t);
  for (int i = 0; i < 3; i++)
    C.t[i] += A.t[i];
  return C;
}
inline SE3 se3_inv(const SE3 &A) {
  SE3 I;
  m3_T(A.R, I.R);
  m3_vec(I.R, A.t, I.t);
  for (int i = 0; i < 3; i++)
    I.t[i] = -I.t[i];
  return I;
}
// quat_ops.h exp_se3: xi = [w, u]
inline SE3 exp_se3(const double *xi) {
  SE3 T;
  double J[9];
  exp_so3(xi, T.R);
  Jl_so3(xi, J);
  m3_vec(J, xi + 3, T.t);
  return T;
}
inline void log_se3(const SE3 &T, double *xi) {
  double J[9], Ji[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  log_so3(T.R, xi);
  Jl_so3(xi, J);
  m3_inv(J, Ji);
  m3_vec(Ji, T.t, xi + 3);
}

// mins_data/BSplineDataset/*.txt: rows [t, px, py, pz, qx, qy, qz, qw] (q = JPL q_GtoI), '#' comments
inline std::vector<std::vector<double>> load_bspline_dataset(const std::string &path) {
  std::ifstream f(path);
  if (!f)
    throw std::runtime_error("cannot open " + path);
  std::vector<std::vector<double>> rows;
  std::string line;
  while (std::getline(f, line)) {
    if (line.empty() || line[0] == '#')
      continue;
    for (auto &ch : line)
      if (ch == ',')
        ch = ' ';
    std::istringstream s(line);
    std::vector<double> v;
    double x;
    while (s >> x)
      v.push_back(x);
    if (v.size() == 8)
      rows.push_back(v);
  }
  return rows;
}

struct BsplineSE3 {
  double dt = 0.05, t_start = 0, t_end = 0;
  std::vector<double> ctrl_t;
  std::vector<SE3> ctrl_T;            // T_ItoG
  std::vector<std::vector<double>> rel; // log(inv(T_i) T_{i+1})

  explicit BsplineSE3(const std::vector<std::vector<double>> &tp) {
    // BsplineSE3.cpp:28-35: control-point spacing = mean spacing of the input, at least 0.05 s
    double sum = 0;
    for (size_t i = 1; i < tp.size(); i++)
      sum += tp[i][0] - tp[i - 1][0];
    dt = std::max(sum / (double)(tp.size() - 1), 0.05);
    std::map<double, SE3> poses;
    for (size_t i = 0; i + 1 < tp.size(); i++) { // :40 (the reference drops the last point)
      SE3 T;
      double R_GtoI[9];
      quat_jpl_to_rot(&tp[i][4], R_GtoI);
      m3_T(R_GtoI, T.R);
      for (int a = 0; a < 3; a++)
        T.t[a] = tp[i][1 + a];
      poses[tp[i][0]] = T;
    }
    std::vector<double> times;
    std::vector<SE3> Ts;
    for (auto &kv : poses) {
      times.push_back(kv.first);
      Ts.push_back(kv.second);
    }
    // :62-88 uniform control points by linear SE(3) interpolation between the bounding input poses
    double t = times[0];
    for (;;) {
      int k = (int)(std::upper_bound(times.begin(), times.end(), t) - times.begin()) - 1;
      if (k < 0 || k + 1 >= (int)times.size())
        break;
      const double lam = (t - times[k]) / (times[k + 1] - times[k]);
      double xi[6];
      log_se3(se3_mul(Ts[k + 1], se3_inv(Ts[k])), xi);
      for (int a = 0; a < 6; a++)
        xi[a] *= lam;
      ctrl_t.push_back(t);
      ctrl_T.push_back(se3_mul(exp_se3(xi), Ts[k]));
      t += dt;
    }
    t_start = times[0] + 2 * dt; // :91
    t_end = ctrl_t.size() > 4 ? ctrl_t[ctrl_t.size() - 3] : ctrl_t.back();
    for (size_t i = 0; i + 1 < ctrl_T.size(); i++) {
      std::vector<double> xi(6);
      log_se3(se3_mul(se3_inv(ctrl_T[i]), ctrl_T[i + 1]), xi.data());
      rel.push_back(xi);
    }
  }
  // T_ItoG at time t (BsplineSE3.cpp:102-136); false outside the spline
  bool pose_T(double t, SE3 *out) const {
    const int i1 = (int)(std::upper_bound(ctrl_t.begin(), ctrl_t.end(), t) - ctrl_t.begin()) - 1;
    if (i1 < 1 || i1 + 2 >= (int)ctrl_t.size())
      return false;
    const double u = (t - ctrl_t[i1]) / (ctrl_t[i1 + 1] - ctrl_t[i1]);
    const double b0 = (5 + 3 * u - 3 * u * u + u * u * u) / 6.0, b1 = (1 + 3 * u + 3 * u * u - 2 * u * u * u) / 6.0, b2 = u * u * u / 6.0;
    double x0[6], x1[6], x2[6];
    for (int a = 0; a < 6; a++)
      x0[a] = b0 * rel[i1 - 1][a], x1[a] = b1 * rel[i1][a], x2[a] = b2 * rel[i1 + 1][a];
    *out = se3_mul(se3_mul(se3_mul(ctrl_T[i1 - 1], exp_se3(x0)), exp_se3(x1)), exp_se3(x2));
    return true;
  }
  bool get_pose(double t, double *R_GtoI, double *p) const {
    SE3 T;
    if (!pose_T(t, &T))
      return false;
    m3_T(T.R, R_GtoI);
    std::memcpy(p, T.t, sizeof(T.t));
    return true;
  }
  // (R_GtoI, p, w_IinI, v_IinG, a_IinG) by 4th-order central differences of the pose
  bool get_kinematics(double t, double *R, double *p, double *w, double *v, double *a, double h = 2e-3) const {
    SE3 T[5];
    for (int k = 0; k < 5; k++)
      if (!pose_T(t + (k - 2) * h, &T[k]))
        return false;
    double Rg[5][9];
    for (int k = 0; k < 5; k++)
      m3_T(T[k].R, Rg[k]);
    for (int i = 0; i < 3; i++) {
      v[i] = (T[0].t[i] - 8 * T[1].t[i] + 8 * T[3].t[i] - T[4].t[i]) / (12 * h);
      a[i] = (-T[0].t[i] + 16 * T[1].t[i] - 30 * T[2].t[i] + 16 * T[3].t[i] - T[4].t[i]) / (12 * h * h);
      p[i] = T[2].t[i];
    }
    // R_GtoI(t+h) = exp_so3(-w h) R_GtoI(t)  (JPL kinematics); Richardson: 4th order
    double Rt[9], M[9], w1[3], w2[3];
    m3_T(Rg[1], Rt);
    m3_mul(Rg[3], Rt, M);
    log_so3(M, w1);
    m3_T(Rg[0], Rt);
    m3_mul(Rg[4], Rt, M);
    log_so3(M, w2);
    for (int i = 0; i < 3; i++)
      w[i] = (4 * (-w1[i] / (2 * h)) - (-w2[i] / (4 * h))) / 3.0;
    std::memcpy(R, Rg[2], sizeof(double) * 9);
    return true;
  }
};

} // namespace mins_sim
