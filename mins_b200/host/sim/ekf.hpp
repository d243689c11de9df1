// Minimal IMU + clone-window EKF core for the closed-loop harness (host C++14, fp64) - SURVEY.md §8(f) N1.
//
// NOT part of the LiDAR hot path: it stands in for the untouched MINS EKF core (mins/src/state/{State,StateHelper,Propagator}),
// which cannot be compiled in this image (no Eigen/Boost/ROS), so that mins_gpu::UpdaterLidar is exercised end to end on a GPU
// and the Monte-Carlo job measures devices, not an interpreter. Restated pieces, with the reference lines they follow:
//   * clone window + marginalisation                     SystemManager.cpp:265-394, StateHelper::marginalize
//   * polynomial pose interpolation and its Jacobians    State.cpp:428-520 (build_polynomial_data), :601-643, :675-826
//   * marginal covariance / EKF update                   StateHelper.cpp:227-254, :121-200
//   * IMU propagation (discrete, first order)            Propagator.cpp (mean: R <- exp(-w dt) R, v, p; Phi, Qd)
// Error-state convention: JPL, R_true = exp_so3(-dth) R_hat (ov_type::JPLQuat::update), positions additive.
// Same simplifications as the Python harness it replaces (mins_b200/sim/ekf.py): polynomial interpolation for the estimate as
// well as the Jacobian (the reference's `use_imu_res: false` branch), constant interpolation-error variances.
#pragma once
#include "se3.hpp"

#include <limits>

namespace mins_sim {

struct Clone {
  double R[9], p[3], R_fej[9], p_fej[3];
  int idx; // first column in the covariance
};

struct PolyData {
  Mat Vinv;
  std::vector<double> c_ori, c_pos;
  std::vector<std::vector<double>> JlInv; // o x 9
  Mat dbth;                               // 3o x 3
  double R0[9], p0[3];
};

// state vector: [IMU 15: th p v bg ba | clones 6 each, oldest first]
struct State {
  double R[9], p[3], v[3], bg[3], ba[3];
  double time = 0;
  int intr_order = 3;
  double g[3] = {0, 0, 9.81};
  std::map<double, Clone> clones;
  double R_ItoL[9], p_IinL[3];
  Mat P;
  double intr_err_ori = 1e-6, intr_err_pos = 1e-5;
  std::map<std::pair<double, int>, PolyData> poly;

  // ---- bookkeeping (State.cpp:936-986) -----------------------------------------------------------------------------
  double oldest_clone_time() const { return clones.empty() ? -std::numeric_limits<double>::infinity() : clones.begin()->first; }
  double newest_clone_time() const { return clones.empty() ? -std::numeric_limits<double>::infinity() : clones.rbegin()->first; }
  double oldest_3rd_clone_time() const {
    if (clones.size() < 3)
      return -std::numeric_limits<double>::infinity();
    auto it = clones.begin();
    ++it, ++it;
    return it->first;
  }
  bool have_clone(double t) const { return clones.find(t) != clones.end(); }

  // ---- IMU propagation ---------------------------------------------------------------------------------------------
  // one IMU interval of length dt with (already averaged) measurements wm, am. noises = (sg, sa, swg, swa)
  void propagate(const double *wm, const double *am, double dt, const double *noises) {
    if (dt <= 0)
      return;
    double w[3], a[3], wdt[3], mwdt[3];
    for (int i = 0; i < 3; i++) {
      w[i] = wm[i] - bg[i], a[i] = am[i] - ba[i];
      wdt[i] = w[i] * dt, mwdt[i] = -wdt[i];
    }
    double dR[9], Jr[9], Rt[9], aG[3], dRJr[9], S[9], tmp[9];
    exp_so3(mwdt, dR);
    Jl_so3(wdt, Jr); // Jr(-w dt) = Jl(w dt)
    m3_T(R, Rt);
    m3_vec(Rt, a, aG);
    for (int i = 0; i < 3; i++)
      aG[i] -= g[i];
    m3_mul(dR, Jr, dRJr);
    Mat F = Mat::eye(15), G(15, 12);
    auto put = [](Mat &M, int r0, int c0, const double *B, double s) {
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
          M(r0 + i, c0 + j) = s * B[3 * i + j];
    };
    double I3[9];
    m3_eye(I3);
    put(F, 0, 0, dR, 1.0);
    put(F, 0, 9, dRJr, -dt);
    skew(a, S);
    m3_mul(Rt, S, tmp);
    put(F, 3, 0, tmp, -0.5 * dt * dt);
    put(F, 3, 6, I3, dt);
    put(F, 3, 12, Rt, -0.5 * dt * dt);
    put(F, 6, 0, tmp, -dt);
    put(F, 6, 12, Rt, -dt);
    put(G, 0, 0, dRJr, -dt);
    put(G, 3, 3, Rt, -0.5 * dt * dt);
    put(G, 6, 3, Rt, -dt);
    put(G, 9, 6, I3, dt);
    put(G, 12, 9, I3, dt);
    Mat Qc(12, 12);
    for (int i = 0; i < 3; i++) {
      Qc(i, i) = noises[0] * noises[0] / dt, Qc(3 + i, 3 + i) = noises[1] * noises[1] / dt;
      Qc(6 + i, 6 + i) = noises[2] * noises[2] / dt, Qc(9 + i, 9 + i) = noises[3] * noises[3] / dt;
    }
    const Mat Qd = mul(mul(G, Qc), transpose(G));
    // mean
    for (int i = 0; i < 3; i++) {
      p[i] += v[i] * dt + 0.5 * aG[i] * dt * dt;
      v[i] += aG[i] * dt;
    }
    m3_mul(dR, R, R);
    // covariance: only the IMU rows/columns change
    const int n = P.r;
    std::vector<double> row(15);
    for (int j = 0; j < n; j++) { // P[:15, :] = F P[:15, :]
      for (int i = 0; i < 15; i++) {
        double s = 0;
        for (int k = 0; k < 15; k++)
          s += F(i, k) * P(k, j);
        row[i] = s;
      }
      for (int i = 0; i < 15; i++)
        P(i, j) = row[i];
    }
    for (int i = 0; i < n; i++) { // P[:, :15] = P[:, :15] F^T
      for (int j = 0; j < 15; j++) {
        double s = 0;
        for (int k = 0; k < 15; k++)
          s += P(i, k) * F(j, k);
        row[j] = s;
      }
      for (int j = 0; j < 15; j++)
        P(i, j) = row[j];
    }
    for (int i = 0; i < 15; i++)
      for (int j = 0; j < 15; j++)
        P(i, j) += Qd(i, j);
    symmetrize();
    time += dt;
  }
  void symmetrize() {
    for (int i = 0; i < P.r; i++)
      for (int j = i + 1; j < P.r; j++) {
        const double m = 0.5 * (P(i, j) + P(j, i));
        P(i, j) = P(j, i) = m;
      }
  }

  // ---- stochastic cloning / marginalisation (StateHelper::augment_clone / marginalize) -----------------------------------
  void augment_clone() {
    const int n = P.r;
    Mat Pn(n + 6, n + 6);
    for (int i = 0; i < n; i++)
      for (int j = 0; j < n; j++)
        Pn(i, j) = P(i, j);
    for (int i = 0; i < 6; i++) // J = [I6 on (th, p)]: the new rows copy the IMU pose rows
      for (int j = 0; j < n; j++) {
        Pn(n + i, j) = P(i, j);
        Pn(j, n + i) = P(j, i);
      }
    for (int i = 0; i < 6; i++)
      for (int j = 0; j < 6; j++)
        Pn(n + i, n + j) = P(i, j);
    P = Pn;
    Clone c;
    std::memcpy(c.R, R, sizeof(R)), std::memcpy(c.p, p, sizeof(p));
    std::memcpy(c.R_fej, R, sizeof(R)), std::memcpy(c.p_fej, p, sizeof(p));
    c.idx = n;
    clones[time] = c;
    poly.clear();
  }
  void marginalize_oldest() {
    const Clone c = clones.begin()->second;
    clones.erase(clones.begin());
    const int n = P.r;
    Mat Pn(n - 6, n - 6);
    for (int i = 0, ii = 0; i < n; i++) {
      if (i >= c.idx && i < c.idx + 6)
        continue;
      for (int j = 0, jj = 0; j < n; j++) {
        if (j >= c.idx && j < c.idx + 6)
          continue;
        Pn(ii, jj++) = P(i, j);
      }
      ii++;
    }
    P = Pn;
    for (auto &kv : clones)
      if (kv.second.idx > c.idx)
        kv.second.idx -= 6;
    poly.clear();
  }

  // ---- polynomial interpolation (State.cpp:428-520, 601-643, 675-826) ----------------------------------------------------
  // intr_order + 1 clone times around t (State::bounding_poses_n): the bracketing pair in the middle when possible
  bool bounding(double t, std::vector<double> &out) const {
    std::vector<double> ts;
    for (auto &kv : clones)
      ts.push_back(kv.first);
    const int n = intr_order + 1;
    if ((int)ts.size() < n || t < ts.front() || t > ts.back())
      return false;
    int k = (int)(std::upper_bound(ts.begin(), ts.end(), t) - ts.begin()) - 1;
    k = std::min(k, (int)ts.size() - 2);
    int start = k - (n / 2 - 1);
    start = std::max(0, std::min(start, (int)ts.size() - n));
    out.assign(ts.begin() + start, ts.begin() + start + n);
    return true;
  }
  const PolyData &poly_data(const std::vector<double> &times, bool fej) {
    const auto key = std::make_pair(times[0], fej ? 1 : 0);
    auto it = poly.find(key);
    if (it != poly.end())
      return it->second;
    const int o = intr_order;
    PolyData d;
    const Clone &c0 = clones.at(times[0]);
    std::memcpy(d.R0, fej ? c0.R_fej : c0.R, sizeof(d.R0));
    std::memcpy(d.p0, fej ? c0.p_fej : c0.p, sizeof(d.p0));
    std::vector<double> d_ori(3 * o), d_pos(3 * o);
    Mat V(3 * o, 3 * o);
    d.dbth = Mat(3 * o, 3);
    double R0t[9];
    m3_T(d.R0, R0t);
    for (int i = 1; i <= o; i++) {
      const Clone &ci = clones.at(times[i]);
      const double *Ri = fej ? ci.R_fej : ci.R, *pi = fej ? ci.p_fej : ci.p;
      double Roi[9], th[3], J[9], Ji[9], JR[9];
      m3_mul(Ri, R0t, Roi);
      log_so3(Roi, th);
      const double dtp = times[i] - times[0];
      for (int a = 0; a < 3; a++) {
        d_ori[3 * (i - 1) + a] = th[a];
        d_pos[3 * (i - 1) + a] = pi[a] - d.p0[a];
        double pw = 1.0;
        for (int j = 1; j <= o; j++) {
          pw *= dtp;
          V(3 * (i - 1) + a, 3 * (j - 1) + a) = pw;
        }
      }
      Jl_so3(th, J);
      m3_inv(J, Ji);
      d.JlInv.push_back(std::vector<double>(Ji, Ji + 9));
      m3_mul(Ji, Roi, JR);
      for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++)
          d.dbth(3 * (i - 1) + a, b) = JR[3 * a + b];
    }
    d.Vinv = inverse(V);
    d.c_ori.assign(3 * o, 0.0), d.c_pos.assign(3 * o, 0.0);
    for (int i = 0; i < 3 * o; i++)
      for (int j = 0; j < 3 * o; j++) {
        d.c_ori[i] += d.Vinv(i, j) * d_ori[j];
        d.c_pos[i] += d.Vinv(i, j) * d_pos[j];
      }
    return poly.emplace(key, d).first->second;
  }
  // State::get_interpolated_pose_poly (State.cpp:601-643)
  bool get_interpolated_pose(double t, double *Rout, double *pout, bool fej = false) {
    auto it = clones.find(t);
    if (it != clones.end()) {
      std::memcpy(Rout, fej ? it->second.R_fej : it->second.R, sizeof(double) * 9);
      std::memcpy(pout, fej ? it->second.p_fej : it->second.p, sizeof(double) * 3);
      return true;
    }
    std::vector<double> times;
    if (!bounding(t, times))
      return false;
    const PolyData &d = poly_data(times, fej);
    const int o = intr_order;
    const double dtm = t - times[0];
    double A_ori[3] = {0, 0, 0}, A_pos[3] = {0, 0, 0}, pw = 1.0;
    for (int i = 0; i < o; i++) {
      pw *= dtm;
      for (int a = 0; a < 3; a++) {
        A_ori[a] += pw * d.c_ori[3 * i + a];
        A_pos[a] += pw * d.c_pos[3 * i + a];
      }
    }
    double E[9];
    exp_so3(A_ori, E);
    m3_mul(E, d.R0, Rout);
    for (int a = 0; a < 3; a++)
      pout[a] = d.p0[a] + A_pos[a];
    return true;
  }
  struct JBlock {
    double t;     // clone time
    double D[36]; // 6 x 6 row-major
  };
  // State::get_interpolated_jacobian (State.cpp:675-826): fej pose and d(pose)/d(clone) blocks
  bool get_interpolated_jacobian(double t, double *Rout, double *pout, std::vector<JBlock> &out) {
    out.clear();
    auto it = clones.find(t);
    if (it != clones.end()) { // exactly at a clone (State::bounding_poses_n returns the clone itself)
      std::memcpy(Rout, it->second.R_fej, sizeof(double) * 9);
      std::memcpy(pout, it->second.p_fej, sizeof(double) * 3);
      JBlock b;
      b.t = t;
      for (int q = 0; q < 36; q++)
        b.D[q] = (q % 7 == 0) ? 1.0 : 0.0;
      out.push_back(b);
      return true;
    }
    std::vector<double> times;
    if (!bounding(t, times))
      return false;
    const PolyData &d = poly_data(times, true);
    const int o = intr_order;
    const double dtm = t - times[0];
    // m_t = [dtm I, dtm^2 I, ...] (3 x 3o); mV = m_t Vinv (3 x 3o)
    std::vector<double> pw(o);
    double A_ori[3] = {0, 0, 0}, A_pos[3] = {0, 0, 0};
    {
      double x = 1.0;
      for (int i = 0; i < o; i++) {
        x *= dtm;
        pw[i] = x;
        for (int a = 0; a < 3; a++) {
          A_ori[a] += x * d.c_ori[3 * i + a];
          A_pos[a] += x * d.c_pos[3 * i + a];
        }
      }
    }
    Mat mV(3, 3 * o);
    for (int a = 0; a < 3; a++)
      for (int j = 0; j < 3 * o; j++) {
        double s = 0;
        for (int i = 0; i < o; i++)
          s += pw[i] * d.Vinv(3 * i + a, j);
        mV(a, j) = s;
      }
    double R0toi[9], Jl[9];
    exp_so3(A_ori, R0toi);
    Jl_so3(A_ori, Jl);
    m3_mul(R0toi, d.R0, Rout);
    for (int a = 0; a < 3; a++)
      pout[a] = d.p0[a] + A_pos[a];
    Mat dth_db(3, 3 * o); // -Jl mV
    for (int a = 0; a < 3; a++)
      for (int j = 0; j < 3 * o; j++)
        dth_db(a, j) = -(Jl[3 * a] * mV(0, j) + Jl[3 * a + 1] * mV(1, j) + Jl[3 * a + 2] * mV(2, j));
    JBlock b0;
    b0.t = times[0];
    std::memset(b0.D, 0, sizeof(b0.D));
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) {
        double s = R0toi[3 * a + b];
        for (int k = 0; k < 3 * o; k++)
          s += dth_db(a, k) * d.dbth(k, b);
        b0.D[6 * a + b] = s;
        double sp = (a == b) ? 1.0 : 0.0; // I + mV dbp, dbp = -[I; I; ...]
        for (int i = 0; i < o; i++)
          sp -= mV(a, 3 * i + b);
        b0.D[6 * (3 + a) + 3 + b] = sp;
      }
    out.push_back(b0);
    for (int i = 0; i < o; i++) {
      JBlock bi;
      bi.t = times[i + 1];
      std::memset(bi.D, 0, sizeof(bi.D));
      for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++) {
          double s = 0;
          for (int k = 0; k < 3; k++)
            s -= dth_db(a, 3 * i + k) * d.JlInv[i][3 * k + b];
          bi.D[6 * a + b] = s;
          bi.D[6 * (3 + a) + 3 + b] = mV(a, 3 * i + b);
        }
      out.push_back(bi);
    }
    return true;
  }

  // ---- covariance access / update (StateHelper.cpp:227-254, 121-200) -----------------------------------------------------
  void marginal_covariance(const std::vector<int> &idx, double *out) const {
    const int n = (int)idx.size();
    for (int i = 0; i < n; i++)
      for (int j = 0; j < n; j++)
        out[(size_t)i * n + j] = P(idx[i], idx[j]);
  }
  // StateHelper::EKFUpdate with R = I; H is rows x idx.size(). Returns false (and changes nothing) if a diagonal would go negative.
  bool ekf_update(const std::vector<int> &idx, const double *H, const double *res, int rows) {
    const int n = P.r, m = (int)idx.size();
    Mat M(n, rows); // P[:, idx] H^T
    for (int i = 0; i < n; i++)
      for (int r = 0; r < rows; r++) {
        double s = 0;
        for (int k = 0; k < m; k++)
          s += P(i, idx[k]) * H[(size_t)r * m + k];
        M(i, r) = s;
      }
    Mat S(rows, rows);
    for (int r = 0; r < rows; r++)
      for (int q = 0; q < rows; q++) {
        double s = (r == q) ? 1.0 : 0.0;
        for (int k = 0; k < m; k++)
          s += H[(size_t)r * m + k] * M(idx[k], q);
        S(r, q) = s;
      }
    for (int r = 0; r < rows; r++)
      for (int q = r + 1; q < rows; q++)
        S(r, q) = S(q, r) = 0.5 * (S(r, q) + S(q, r));
    Mat Kt; // S^-1 M^T (rows x n)
    if (!chol_solve(S, transpose(M), Kt))
      return false;
    Mat Pn = P;
    for (int i = 0; i < n; i++)
      for (int j = 0; j < n; j++) {
        double s = 0;
        for (int r = 0; r < rows; r++)
          s += Kt(r, i) * M(j, r);
        Pn(i, j) -= s;
      }
    for (int i = 0; i < n; i++)
      if (Pn(i, i) < 0)
        return false;
    P = Pn;
    symmetrize();
    std::vector<double> dx(n, 0.0);
    for (int i = 0; i < n; i++)
      for (int r = 0; r < rows; r++)
        dx[i] += Kt(r, i) * res[r];
    apply(dx.data());
    return true;
  }
  void apply(const double *dx) {
    double m[3], E[9];
    for (int i = 0; i < 3; i++)
      m[i] = -dx[i];
    exp_so3(m, E);
    m3_mul(E, R, R);
    for (int i = 0; i < 3; i++) {
      p[i] += dx[3 + i], v[i] += dx[6 + i], bg[i] += dx[9 + i], ba[i] += dx[12 + i];
    }
    for (auto &kv : clones) {
      Clone &c = kv.second;
      for (int i = 0; i < 3; i++)
        m[i] = -dx[c.idx + i];
      exp_so3(m, E);
      m3_mul(E, c.R, c.R);
      for (int i = 0; i < 3; i++)
        c.p[i] += dx[c.idx + 3 + i];
    }
    for (auto it = poly.begin(); it != poly.end();) // fej polynomials stay valid
      it = it->first.second ? std::next(it) : poly.erase(it);
  }
};

} // namespace mins_sim
