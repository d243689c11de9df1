// =====================================================================================================
// TEST INFRASTRUCTURE ONLY — CPU oracle for the MINS LiDAR measurement-update path.
// Nothing under oracle/ is linked, imported or executed by the product (mins_b200/, include/): only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it, as the checker / baseline.
//
// This file is a plain-C++ (no Eigen/PCL/Boost — none exist in this image) RESTATEMENT of the reference's
// algorithm, function by function, each citing the reference file:line it follows (paths relative to
// /root/reference). The kNN/map part of the path is additionally available as the reference's own compiled
// code (oracle/ref_ikd.cpp -> oracle/_ref/libikd_ref.so) and this restatement is pinned against it and against
// libnabo's scan fixtures in tests/ (see DESIGN.md "Oracle pinning"). Third-party arithmetic whose source is
// NOT in the reference tree (PCL 1.10 VoxelGrid / transformPointCloud, Eigen 3.3.7 JacobiSVD /
// ColPivHouseholderQR / makeGivens, Boost chi-squared quantile) is restated from the published algorithms:
// for those pieces parity is "unpinned" (no reference test or binary exists to check against).
// =====================================================================================================
#include "../include/mins_lidar.h"
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <map>
#include <thread>
#include <vector>

namespace {

// ---------------------------------------------------------------------------------------------
// tiny dense helpers (row-major, double)
// ---------------------------------------------------------------------------------------------
struct Mat {
  int r = 0, c = 0;
  std::vector<double> a;
  Mat() {}
  Mat(int r_, int c_) : r(r_), c(c_), a((size_t)r_ * c_, 0.0) {}
  double &operator()(int i, int j) { return a[(size_t)i * c + j]; }
  double operator()(int i, int j) const { return a[(size_t)i * c + j]; }
};

inline void mat3_mul(const double *A, const double *B, double *C) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
inline void mat3_vec(const double *A, const double *v, double *o) {
  for (int i = 0; i < 3; i++)
    o[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
}
// ov_core quat_ops.h:135 skew_x
inline void skew(const double *w, double *S) {
  S[0] = 0, S[1] = -w[2], S[2] = w[1];
  S[3] = w[2], S[4] = 0, S[5] = -w[0];
  S[6] = -w[1], S[7] = w[0], S[8] = 0;
}

// Singular values of an m x n (m >= n) matrix by one-sided Jacobi (Hestenes). Stands in for
// Eigen::JacobiSVD<MatrixXd> at LidarHelper.cpp:305-309 (any backward-stable SVD returns the same values to ~1e-15).
void singular_values(const Mat &Ain, std::vector<double> &sv) {
  Mat A = Ain;
  int m = A.r, n = A.c;
  for (int sweep = 0; sweep < 60; sweep++) {
    double off = 0;
    for (int p = 0; p < n - 1; p++)
      for (int q = p + 1; q < n; q++) {
        double alpha = 0, beta = 0, gamma = 0;
        for (int i = 0; i < m; i++) {
          alpha += A(i, p) * A(i, p);
          beta += A(i, q) * A(i, q);
          gamma += A(i, p) * A(i, q);
        }
        if (gamma == 0.0)
          continue;
        off = std::max(off, std::fabs(gamma) / std::sqrt(alpha * beta));
        double zeta = (beta - alpha) / (2.0 * gamma);
        double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
        double cs = 1.0 / std::sqrt(1.0 + t * t), sn = cs * t;
        for (int i = 0; i < m; i++) {
          double x = A(i, p), y = A(i, q);
          A(i, p) = cs * x - sn * y;
          A(i, q) = sn * x + cs * y;
        }
      }
    if (off < 1e-15)
      break;
  }
  sv.assign(n, 0.0);
  for (int j = 0; j < n; j++) {
    double s = 0;
    for (int i = 0; i < m; i++)
      s += A(i, j) * A(i, j);
    sv[j] = std::sqrt(s);
  }
  std::sort(sv.begin(), sv.end(), std::greater<double>());
}

// x = argmin ||A x - b|| by Householder QR with column pivoting (largest remaining column norm first), the
// algorithm of Eigen::ColPivHouseholderQR::solve used at LidarHelper.cpp:315. Rank decision as Eigen:
// pivots below eps * n * max_pivot are treated as zero.
void colpiv_qr_solve(const Mat &Ain, const std::vector<double> &bin, std::vector<double> &x) {
  Mat A = Ain;
  int m = A.r, n = A.c;
  std::vector<double> b = bin;
  std::vector<int> perm(n);
  for (int j = 0; j < n; j++)
    perm[j] = j;
  std::vector<double> diag(n, 0.0);
  double maxpivot = 0;
  int rank = n;
  for (int k = 0; k < n; k++) {
    // pivot: column with the largest remaining norm
    int best = k;
    double bestn = -1;
    for (int j = k; j < n; j++) {
      double s = 0;
      for (int i = k; i < m; i++)
        s += A(i, j) * A(i, j);
      if (s > bestn)
        bestn = s, best = j;
    }
    if (best != k) {
      for (int i = 0; i < m; i++)
        std::swap(A(i, k), A(i, best));
      std::swap(perm[k], perm[best]);
    }
    // Householder vector for column k
    double normx = std::sqrt(bestn);
    if (normx == 0.0) {
      rank = k;
      break;
    }
    double alpha = A(k, k) > 0 ? -normx : normx;
    std::vector<double> v(m, 0.0);
    v[k] = A(k, k) - alpha;
    for (int i = k + 1; i < m; i++)
      v[i] = A(i, k);
    double vnorm2 = 0;
    for (int i = k; i < m; i++)
      vnorm2 += v[i] * v[i];
    if (vnorm2 > 0) {
      for (int j = k; j < n; j++) {
        double dot = 0;
        for (int i = k; i < m; i++)
          dot += v[i] * A(i, j);
        double f = 2.0 * dot / vnorm2;
        for (int i = k; i < m; i++)
          A(i, j) -= f * v[i];
      }
      double dot = 0;
      for (int i = k; i < m; i++)
        dot += v[i] * b[i];
      double f = 2.0 * dot / vnorm2;
      for (int i = k; i < m; i++)
        b[i] -= f * v[i];
    }
    diag[k] = A(k, k);
    maxpivot = std::max(maxpivot, std::fabs(diag[k]));
  }
  double thr = std::numeric_limits<double>::epsilon() * (double)std::min(m, n) * maxpivot;
  int nz = 0;
  for (int k = 0; k < rank; k++)
    if (std::fabs(diag[k]) > thr)
      nz++;
  std::vector<double> y(n, 0.0);
  for (int k = nz - 1; k >= 0; k--) {
    double s = b[k];
    for (int j = k + 1; j < nz; j++)
      s -= A(k, j) * y[j];
    y[k] = s / A(k, k);
  }
  x.assign(n, 0.0);
  for (int k = 0; k < n; k++)
    x[perm[k]] = y[k];
}

// Eigen::JacobiRotation<double>::makeGivens(p, q) (real case): G = [c s; -s c] with G^* [p;q] = [r;0].
inline void make_givens(double p, double q, double &c, double &s) {
  if (q == 0.0) {
    c = p < 0 ? -1.0 : 1.0;
    s = 0.0;
  } else if (p == 0.0) {
    c = 0.0;
    s = q < 0 ? 1.0 : -1.0;
  } else if (std::fabs(p) > std::fabs(q)) {
    double t = q / p, u = std::sqrt(1.0 + t * t);
    if (p < 0)
      u = -u;
    c = 1.0 / u;
    s = -t * c;
  } else {
    double t = p / q, u = std::sqrt(1.0 + t * t);
    if (q < 0)
      u = -u;
    s = -1.0 / u;
    c = -t * s;
  }
}
// rows (i-1, i) <- G^* rows : x' = c x - s y ; y' = s x + c y
inline void rot_rows(Mat &M, int i0, int i1, int c0, double c, double s) {
  for (int j = c0; j < M.c; j++) {
    double x = M(i0, j), y = M(i1, j);
    M(i0, j) = c * x - s * y;
    M(i1, j) = s * x + c * y;
  }
}

// StateHelper::Givens_Rotation(A, B, C) — StateHelper.cpp:660-680
void givens_ABC(Mat &A, Mat &B, std::vector<double> &C) {
  for (int n = 0; n < A.c; ++n)
    for (int m = A.r - 1; m > n; m--) {
      if (A(m, n) == 0.0)
        continue;
      double c, s;
      make_givens(A(m - 1, n), A(m, n), c, s);
      rot_rows(A, m - 1, m, n, c, s);
      A(m, n) = 0;
      rot_rows(B, m - 1, m, 0, c, s);
      double x = C[m - 1], y = C[m];
      C[m - 1] = c * x - s * y;
      C[m] = s * x + c * y;
    }
}
// StateHelper::Givens_Rotation(A, B) — StateHelper.cpp:682-701
void givens_AB(Mat &A, std::vector<double> &B) {
  for (int n = 0; n < A.c; ++n)
    for (int m = A.r - 1; m > n; m--) {
      if (A(m, n) == 0.0)
        continue;
      double c, s;
      make_givens(A(m - 1, n), A(m, n), c, s);
      rot_rows(A, m - 1, m, n, c, s);
      A(m, n) = 0;
      double x = B[m - 1], y = B[m];
      B[m - 1] = c * x - s * y;
      B[m] = s * x + c * y;
    }
}

// general small inverse by Gauss-Jordan with partial pivoting (stands in for Eigen's PartialPivLU inverse)
bool invert(Mat &M) {
  int n = M.r;
  Mat I(n, n);
  for (int i = 0; i < n; i++)
    I(i, i) = 1;
  for (int k = 0; k < n; k++) {
    int p = k;
    for (int i = k + 1; i < n; i++)
      if (std::fabs(M(i, k)) > std::fabs(M(p, k)))
        p = i;
    if (M(p, k) == 0.0)
      return false;
    if (p != k)
      for (int j = 0; j < n; j++)
        std::swap(M(k, j), M(p, j)), std::swap(I(k, j), I(p, j));
    double d = 1.0 / M(k, k);
    for (int j = 0; j < n; j++)
      M(k, j) *= d, I(k, j) *= d;
    for (int i = 0; i < n; i++)
      if (i != k) {
        double f = M(i, k);
        if (f != 0.0)
          for (int j = 0; j < n; j++)
            M(i, j) -= f * M(k, j), I(i, j) -= f * I(k, j);
      }
  }
  M = I;
  return true;
}

// KD_TREE::calc_dist — ikd_Tree.cpp:1406-1410; fp32, no FMA contraction (reference builds for baseline x86-64,
// mins/CMakeLists.txt:15). This file is compiled with -ffp-contract=off.
inline float calc_dist(const float *a, const float *b) {
  float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
  float d = dx * dx + dy * dy;
  d = d + dz * dz;
  return d;
}

} // namespace

extern "C" {

// ---------------------------------------------------------------------------------------------------------
// A4  exact k-NN, brute force. Semantics = libnabo BruteForceSearch::knn (thirdparty/libnabo/nabo/
// brute_force_cpu.cpp:71-100): scan candidates in ascending index, replace the worst kept only if strictly
// closer => total order (d2, index) ascending; distance arithmetic = KD_TREE::calc_dist. idx = -1 / d2 = inf
// where fewer than k candidates have d2 <= max_d2. ok = LidarHelper::get_neighbors' verdict (LidarHelper.cpp:280-285).
// ---------------------------------------------------------------------------------------------------------
void orc_knn_brute(const float *map_xyzi, int M, const float *q_xyzi, int N, int k, float max_d2, int32_t *idx,
                   float *d2, uint8_t *ok, int threads) {
  if (threads < 1)
    threads = 1;
  auto work = [&](int s, int e) {
    std::vector<float> bd(k);
    std::vector<int32_t> bi(k);
    for (int i = s; i < e; i++) {
      for (int j = 0; j < k; j++)
        bd[j] = INFINITY, bi[j] = -1;
      const float *q = q_xyzi + 4 * (size_t)i;
      for (int m = 0; m < M; m++) {
        float d = calc_dist(map_xyzi + 4 * (size_t)m, q);
        if (d <= max_d2 && d < bd[k - 1]) {
          int p = k - 1;
          while (p > 0 && d < bd[p - 1]) {
            bd[p] = bd[p - 1];
            bi[p] = bi[p - 1];
            p--;
          }
          bd[p] = d;
          bi[p] = m;
        }
      }
      for (int j = 0; j < k; j++) {
        if (idx)
          idx[(size_t)i * k + j] = bi[j];
        if (d2)
          d2[(size_t)i * k + j] = bd[j];
      }
      if (ok)
        ok[i] = (bi[k - 1] >= 0 && !((double)bd[k - 1] > (double)max_d2)) ? 1 : 0;
    }
  };
  std::vector<std::thread> th;
  int chunk = (N + threads - 1) / threads;
  for (int t = 0; t < threads; t++) {
    int s = t * chunk, e = std::min(N, s + chunk);
    if (s < e)
      th.emplace_back(work, s, e);
  }
  for (auto &t : th)
    t.join();
}

// ---------------------------------------------------------------------------------------------------------
// boost::math::quantile(chi_squared(dof), 0.95) — UpdaterStatistics.cpp:31-33 fills chi_squared_table[i] with it and
// get_chi2 (:110-111) reads chi_squared_table[res.rows()]. Computed here (not tabulated): regularised lower incomplete
// gamma P(dof/2, x/2) by its power series, root by bisection to the last bit. tests/test_oracle_math.py pins it to
// scipy.stats.chi2.ppf.
// ---------------------------------------------------------------------------------------------------------
static double reg_lower_gamma(double a, double x) {
  if (x <= 0)
    return 0.0;
  double term = 1.0 / a, sum = term;
  for (int n = 1; n < 2000; n++) {
    term *= x / (a + n);
    sum += term;
    if (term < sum * 1e-18)
      break;
  }
  return sum * std::exp(-x + a * std::log(x) - std::lgamma(a));
}
static double chi2_quantile95_solve(int dof) {
  double lo = 0.0, hi = 10.0 * dof + 50.0;
  for (int it = 0; it < 200; it++) {
    const double mid = 0.5 * (lo + hi);
    if (reg_lower_gamma(0.5 * dof, 0.5 * mid) < 0.95)
      lo = mid;
    else
      hi = mid;
  }
  return 0.5 * (lo + hi);
}
// the table UpdaterStatistics' constructor builds (max_chi_size entries there; 64 cover every row count of this path)
double orc_chi2_quantile95(int dof) {
  struct Table {
    double q[64];
    Table() {
      q[0] = 0.0;
      for (int i = 1; i < 64; i++)
        q[i] = chi2_quantile95_solve(i);
    }
  };
  static const Table t; // thread-safe initialisation (C++11)
  return dof < 64 ? t.q[dof < 0 ? 0 : dof] : chi2_quantile95_solve(dof);
}

// ---------------------------------------------------------------------------------------------------------
// A5  LidarHelper::compute_plane — LidarHelper.cpp:291-328. nb: k x float[4] neighbour points (map frame).
// returns 1 on success (plane = [abc, d], un-normalised), 0 on rejection. cond (may be NULL) = sigma_max/sigma_min.
// ---------------------------------------------------------------------------------------------------------
int orc_compute_plane(const float *nb_xyzi, int k, double plane_max_condi, double plane_max_p2pd, double *plane,
                      double *cond) {
  Mat A(k, 3);
  double d = 0;
  for (int j = 0; j < k; j++) { // :295-300
    A(j, 0) = nb_xyzi[4 * j + 0];
    A(j, 1) = nb_xyzi[4 * j + 1];
    A(j, 2) = nb_xyzi[4 * j + 2];
    d += std::sqrt(A(j, 0) * A(j, 0) + A(j, 1) * A(j, 1) + A(j, 2) * A(j, 2));
  }
  d /= k; // :301
  std::vector<double> b(k, -d), sv, abc; // :302
  singular_values(A, sv);                // :305-308
  double condA = sv[0] / sv[2];          // :309
  if (cond)
    *cond = condA;
  if (condA > plane_max_condi) // :310
    return 0;
  colpiv_qr_solve(A, b, abc); // :315
  plane[0] = abc[0], plane[1] = abc[1], plane[2] = abc[2], plane[3] = d;
  double nrm = std::sqrt(abc[0] * abc[0] + abc[1] * abc[1] + abc[2] * abc[2]);
  double pl[4] = {plane[0] / nrm, plane[1] / nrm, plane[2] / nrm, plane[3] / nrm}; // :319
  for (int j = 0; j < k; j++) {                                                    // :320-325
    const float *pt = nb_xyzi + 4 * j;
    if (std::fabs(pl[0] * pt[0] + pl[1] * pt[1] + pl[2] * pt[2] + pl[3]) > plane_max_p2pd)
      return 0;
  }
  return 1;
}

// ---------------------------------------------------------------------------------------------------------
// A6-A8  one scan point: body of the thread loop, UpdaterLidar.cpp:279-368.
//   pfinL_f  : the point in the LiDAR frame (float xyz, `lidar->p(i)` :281)
//   nb_xyzi  : its k neighbours in the map (result of get_neighbors :287, already verified ok)
//   Hp (3 x total_hx, row-major) and rp (3): the projected, whitened rows kept in vec_lin[i] (:366-367)
// returns ml_point_row.status (0 accepted, 2 plane fail, 3 bad match, 4 chi2 reject). row (may be NULL) gets the
// invariants; hrow (may be NULL, total_hx) the whitened last row of H before projection (:355).
// ---------------------------------------------------------------------------------------------------------
int orc_point(const ml_scan_consts *c, const double *P, const float *pfinL_f, const float *nb_xyzi, double *Hp,
              double *rp, ml_point_row *row, double *hrow) {
  const int k = c->map_ngbr_num, hx = c->total_hx;
  double pfinL[3] = {pfinL_f[0], pfinL_f[1], pfinL_f[2]}; // :281
  double pfinM_est[3], pfinM_fej[3], t[3];
  mat3_vec(c->RLtoM_est, pfinL, t); // :282
  for (int i = 0; i < 3; i++)
    pfinM_est[i] = c->pLinM_est[i] + t[i];
  mat3_vec(c->RLtoM_fej, pfinL, t); // :283
  for (int i = 0; i < 3; i++)
    pfinM_fej[i] = c->pLinM_fej[i] + t[i];
  if (row) {
    memset(row, 0, sizeof(*row));
  }

  double plane[4];
  if (!orc_compute_plane(nb_xyzi, k, c->plane_max_condi, c->plane_max_p2pd, plane, nullptr)) { // :292
    if (row)
      row->status = 2;
    return 2;
  }
  if (row)
    memcpy(row->plane, plane, sizeof(plane));
  // :296 discard bad match (one-sided, un-normalised)
  if (-(plane[0] * pfinM_est[0] + plane[1] * pfinM_est[1] + plane[2] * pfinM_est[2]) - plane[3] > c->plane_max_p2pd * 3) {
    if (row)
      row->status = 3;
    return 3;
  }
  const int rows = k + 1; // :301
  std::vector<double> res(rows, 0.0);
  Mat Hf(rows, 3);
  for (int j = 0; j < k; j++) { // :305-310
    double pn[3] = {nb_xyzi[4 * j], nb_xyzi[4 * j + 1], nb_xyzi[4 * j + 2]};
    res[j] = -(plane[0] * pn[0] + plane[1] * pn[1] + plane[2] * pn[2]) - plane[3];
    for (int a = 0; a < 3; a++)
      Hf(j, a) = pn[a];
  }
  res[rows - 1] = -(plane[0] * pfinM_est[0] + plane[1] * pfinM_est[1] + plane[2] * pfinM_est[2]) - plane[3]; // :313
  for (int a = 0; a < 3; a++)
    Hf(rows - 1, a) = pfinM_fej[a]; // :314

  // Jacobians :320-329
  const double *n = plane; // dz_dpfinM_bottom_row
  double dI[18], dA[18], dC[18], S[9], M1[9], M2[9], v[3], w[3];
  // dpfinM_dI = [-RGtoM*RItoG*skew(RLtoI*(pfinL - pIinL)), RGtoM]
  for (int i = 0; i < 3; i++)
    v[i] = pfinL[i] - c->pIinL_fej[i];
  mat3_vec(c->RLtoI_fej, v, w);
  skew(w, S);
  mat3_mul(c->RGtoM_fej, c->RItoG_fej, M1);
  mat3_mul(M1, S, M2);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      dI[6 * i + j] = -M2[3 * i + j];
      dI[6 * i + 3 + j] = c->RGtoM_fej[3 * i + j];
    }
  // dpfinM_dA = [RItoL*skew(RGtoA*(pIinG + RItoG*pLinI - pAinG + RLtoG*pfinL)), -RGtoM]
  double a1[3], a2[3];
  mat3_vec(c->RItoG_fej, c->pLinI_fej, a1);
  mat3_vec(c->RLtoG_fej, pfinL, a2);
  for (int i = 0; i < 3; i++)
    v[i] = c->pIinG_fej[i] + a1[i] - c->pAinG_fej[i] + a2[i];
  mat3_vec(c->RGtoA_fej, v, w);
  skew(w, S);
  mat3_mul(c->RItoL_fej, S, M1);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      dA[6 * i + j] = M1[3 * i + j];
      dA[6 * i + 3 + j] = -c->RGtoM_fej[3 * i + j];
    }
  // dpfinM_dcalib = [skew(RGtoM*(pIinG-pAinG)) - skew(RLtoM*(pIinL-pfinL)) + RLtoM*skew(pIinL-pfinL), I - RLtoM]
  double S1[9], S2[9], S3[9];
  for (int i = 0; i < 3; i++)
    v[i] = c->pIinG_fej[i] - c->pAinG_fej[i];
  mat3_vec(c->RGtoM_fej, v, w);
  skew(w, S1);
  for (int i = 0; i < 3; i++)
    v[i] = c->pIinL_fej[i] - pfinL[i];
  mat3_vec(c->RLtoM_fej, v, w);
  skew(w, S2);
  skew(v, S3);
  mat3_mul(c->RLtoM_fej, S3, M1);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      dC[6 * i + j] = S1[3 * i + j] - S2[3 * i + j] + M1[3 * i + j];
      dC[6 * i + 3 + j] = (i == j ? 1.0 : 0.0) - c->RLtoM_fej[3 * i + j];
    }
  // n * dpfinM_d*
  double nI[6], nA[6], nC[6];
  for (int j = 0; j < 6; j++) {
    nI[j] = n[0] * dI[j] + n[1] * dI[6 + j] + n[2] * dI[12 + j];
    nA[j] = n[0] * dA[j] + n[1] * dA[6 + j] + n[2] * dA[12 + j];
    nC[j] = n[0] * dC[j] + n[1] * dC[6 + j] + n[2] * dC[12 + j];
  }
  // :332-337 interpolation-error inflation (uses dpfinM_dI only — reference quirk kept)
  double intr_std = 0.0;
  if (!c->at_clone_m || !c->at_clone_a)
    intr_std += std::sqrt(c->ie_o * (nI[0] * nI[0] + nI[1] * nI[1] + nI[2] * nI[2]) +
                          c->ie_p * (nI[3] * nI[3] + nI[4] * nI[4] + nI[5] * nI[5]));
  // :342-350
  Mat H(rows, hx);
  for (int b = 0; b < c->nI; b++) {
    const ml_jac_block &B = c->dTdxI[b];
    for (int cc = 0; cc < B.width; cc++) {
      double s = 0;
      for (int r = 0; r < 6; r++)
        s += nI[r] * B.D[6 * r + cc];
      H(rows - 1, B.col + cc) += s;
    }
  }
  for (int b = 0; b < c->nA; b++) {
    const ml_jac_block &B = c->dTdxA[b];
    for (int cc = 0; cc < B.width; cc++) {
      double s = 0;
      for (int r = 0; r < 6; r++)
        s += nA[r] * B.D[6 * r + cc];
      H(rows - 1, B.col + cc) += s;
    }
  }
  if (c->calib_col >= 0)
    for (int cc = 0; cc < 6; cc++)
      H(rows - 1, c->calib_col + cc) += nC[cc];
  // :352-359 whitening
  double nn = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
  double raw_noise = nn * c->raw_noise + intr_std;
  double map_noise = nn * c->map_noise + intr_std;
  for (int j = 0; j < hx; j++)
    H(rows - 1, j) /= raw_noise;
  for (int a = 0; a < 3; a++)
    Hf(rows - 1, a) /= raw_noise;
  for (int j = 0; j < k; j++)
    for (int a = 0; a < 3; a++)
      Hf(j, a) /= map_noise;
  res[rows - 1] /= raw_noise;
  for (int j = 0; j < k; j++)
    res[j] /= map_noise;
  if (hrow)
    for (int j = 0; j < hx; j++)
      hrow[j] = H(rows - 1, j);
  double res_last = res[rows - 1];

  // :362 nullspace projection — StateHelper.cpp:645-658
  givens_ABC(Hf, H, res);
  const int pr = rows - 3; // rows kept
  Mat Hn(pr, hx);
  std::vector<double> rn(pr);
  for (int i = 0; i < pr; i++) {
    rn[i] = res[3 + i];
    for (int j = 0; j < hx; j++)
      Hn(i, j) = H(3 + i, j);
  }
  // :365 Chi2Check -> get_chi2, UpdaterStatistics.cpp:94-117 with R = I
  Mat HP(pr, hx), Sm(pr, pr);
  for (int i = 0; i < pr; i++)
    for (int j = 0; j < hx; j++) {
      double s = 0;
      for (int l = 0; l < hx; l++)
        s += Hn(i, l) * P[(size_t)l * hx + j];
      HP(i, j) = s;
    }
  for (int i = 0; i < pr; i++)
    for (int j = 0; j < pr; j++) {
      double s = 0;
      for (int l = 0; l < hx; l++)
        s += HP(i, l) * Hn(j, l);
      Sm(i, j) = s + (i == j ? 1.0 : 0.0);
    }
  for (int i = 0; i < pr; i++) // selfadjointView<Upper>
    for (int j = 0; j < i; j++)
      Sm(i, j) = Sm(j, i);
  double chi = std::numeric_limits<double>::quiet_NaN();
  if (invert(Sm)) {
    chi = 0;
    for (int i = 0; i < pr; i++)
      for (int j = 0; j < pr; j++)
        chi += rn[i] * Sm(i <= j ? i : j, i <= j ? j : i) * rn[j];
  }
  if (row) {
    row->res = res_last;
    row->chi2 = chi;
    row->raw_n = raw_noise;
    // invariants of the projected rows: H'^T H' = w h h^T, H'^T r' = h vr  (H' = v h^T)
    // recover v from any non-zero column of H': use least squares against hrow
    double hh = 0;
    std::vector<double> hv(hx);
    // h (whitened) is only available if hrow was requested; recompute from H' : H'^T H' trace = w * ||h||^2
    double tr = 0;
    for (int i = 0; i < pr; i++)
      for (int j = 0; j < hx; j++)
        tr += Hn(i, j) * Hn(i, j);
    if (hrow) {
      for (int j = 0; j < hx; j++)
        hh += hrow[j] * hrow[j];
      row->w = hh > 0 ? tr / hh : 0;
      double s = 0; // h^T (H'^T r') / ||h||^2
      for (int j = 0; j < hx; j++) {
        double g = 0;
        for (int i = 0; i < pr; i++)
          g += Hn(i, j) * rn[i];
        s += hrow[j] * g;
      }
      row->vr = hh > 0 ? s / hh : 0;
    }
  }
  const double thr = c->chi2_mult * orc_chi2_quantile95(pr); // UpdaterStatistics.cpp:110-111: chi_squared_table[res.rows()]
  int status = (std::isnan(chi) || !(chi < thr)) ? 4 : 0; // UpdaterStatistics.cpp:117,59
  if (row)
    row->status = status;
  if (Hp)
    for (int i = 0; i < pr; i++)
      for (int j = 0; j < hx; j++)
        Hp[(size_t)i * hx + j] = Hn(i, j);
  if (rp)
    for (int i = 0; i < pr; i++)
      rp[i] = rn[i];
  return status;
}

// ---------------------------------------------------------------------------------------------------------
// A4-A9 whole scan: UpdaterLidar.cpp:266-397. Inputs: scan points in L (pointcloud, float4), the same points in
// the map frame (pointcloud_in_map, float4; the kNN query, :287), the map (flatten order). kNN by brute force
// unless knn_idx is given. Builds vec_lin, stacks H_big/res_big (:376-394) and returns
//   HtH = H_big^T H_big, Htr = H_big^T res_big   (what the GPU path returns),
//   optionally Hc/rc = the Givens-compressed system of :397 (Hc: total_hx x total_hx, rc: total_hx; n_c rows valid).
// Threading = the reference's static chunks (:269-277).
// ---------------------------------------------------------------------------------------------------------
static int scan_linsys_core(const ml_scan_consts *c, const double *P, const float *map_xyzi, int M, const float *scan_L,
                    const float *scan_M, int N, const int32_t *knn_idx_in, const float *nb_in, const uint8_t *nb_ok_in,
                    double *HtH, double *Htr, ml_counters *cnt,
                    ml_point_row *rows_out, double *h_rows_out, double *Hc, double *rc, int *n_c, int threads) {
  const int k = c->map_ngbr_num, hx = c->total_hx;
  std::vector<int32_t> idx((size_t)N * k);
  std::vector<uint8_t> ok(N);
  std::vector<float> d2((size_t)N * k);
  if (threads < 1)
    threads = 1;
  if (nb_in) {
    for (int i = 0; i < N; i++)
      ok[i] = nb_ok_in[i];
  } else if (knn_idx_in) {
    memcpy(idx.data(), knn_idx_in, sizeof(int32_t) * (size_t)N * k);
    for (int i = 0; i < N; i++) {
      bool good = true;
      for (int j = 0; j < k; j++)
        good = good && idx[(size_t)i * k + j] >= 0;
      if (good) {
        float dk = calc_dist(map_xyzi + 4 * (size_t)idx[(size_t)i * k + k - 1], scan_M + 4 * (size_t)i);
        good = !((double)dk > c->map_ngbr_max_d);
      }
      ok[i] = good;
    }
  } else {
    orc_knn_brute(map_xyzi, M, scan_M, N, k, (float)c->map_ngbr_max_d, idx.data(), d2.data(), ok.data(), threads);
  }
  const int pr = k + 1 - 3;
  std::vector<int> status(N, 1);
  std::vector<double> Hall((size_t)N * pr * hx), rall((size_t)N * pr);
  std::vector<ml_point_row> rows(N);
  // the per-point Jacobian rows and the (w, vr) invariants are test exports: not formed on the timed baseline path
  const bool want_rows = rows_out || h_rows_out;
  std::vector<double> hrows(want_rows ? (size_t)N * hx : 0, 0.0);
  auto work = [&](int s, int e) {
    std::vector<float> nb(4 * (size_t)k);
    for (int i = s; i < e; i++) {
      memset(&rows[i], 0, sizeof(ml_point_row));
      rows[i].status = 1;
      if (!ok[i])
        continue;
      for (int j = 0; j < k; j++)
        memcpy(&nb[4 * j], nb_in ? nb_in + 4 * ((size_t)i * k + j) : map_xyzi + 4 * (size_t)idx[(size_t)i * k + j], 4 * sizeof(float));
      status[i] = orc_point(c, P, scan_L + 4 * (size_t)i, nb.data(), &Hall[(size_t)i * pr * hx], &rall[(size_t)i * pr],
                            &rows[i], want_rows ? &hrows[(size_t)i * hx] : nullptr);
    }
  };
  {
    std::vector<std::thread> th;
    int chunk = N / threads;
    for (int t = 0; t < threads; t++) {
      int s = t * chunk, e = (t == threads - 1) ? N : (t + 1) * chunk;
      th.emplace_back(work, s, e);
    }
    for (auto &t : th)
      t.join();
  }
  ml_counters cn;
  memset(&cn, 0, sizeof(cn));
  cn.n_points = N;
  int total_res = 0;
  for (int i = 0; i < N; i++) {
    switch (status[i]) {
    case 0: cn.n_acc++; total_res += pr; break;
    case 1: cn.n_knn_fail++; break;
    case 2: cn.n_plane_fail++; break;
    case 3: cn.n_badmatch++; break;
    default: cn.n_chi_rej++; break;
    }
    if ((status[i] == 0 || status[i] == 4) && !std::isnan(rows[i].chi2)) {
      cn.sum_chi += rows[i].chi2;
      cn.sum_chi2 += rows[i].chi2 * rows[i].chi2;
      for (int a = 0; a < pr; a++) {
        double r = rall[(size_t)i * pr + a];
        cn.sum_res += r;
        cn.sum_res2 += r * r;
      }
    }
  }
  if (cnt)
    *cnt = cn;
  if (rows_out)
    memcpy(rows_out, rows.data(), sizeof(ml_point_row) * (size_t)N);
  if (h_rows_out)
    memcpy(h_rows_out, hrows.data(), sizeof(double) * (size_t)N * hx);
  // :384-394 stack
  Mat Hbig(total_res, hx);
  std::vector<double> rbig(total_res);
  int r0 = 0;
  for (int i = 0; i < N; i++) {
    if (status[i] != 0)
      continue;
    for (int a = 0; a < pr; a++) {
      rbig[r0] = rall[(size_t)i * pr + a];
      memcpy(&Hbig.a[(size_t)r0 * hx], &Hall[((size_t)i * pr + a) * hx], sizeof(double) * hx);
      r0++;
    }
  }
  if (HtH) {
    for (int a = 0; a < hx * hx; a++)
      HtH[a] = 0;
    for (int r = 0; r < total_res; r++) {
      const double *hr = &Hbig.a[(size_t)r * hx];
      for (int a = 0; a < hx; a++) {
        double ha = hr[a];
        if (ha == 0.0)
          continue;
        for (int b = 0; b < hx; b++)
          HtH[a * hx + b] += ha * hr[b];
      }
    }
  }
  if (Htr) {
    for (int a = 0; a < hx; a++)
      Htr[a] = 0;
    for (int r = 0; r < total_res; r++)
      for (int a = 0; a < hx; a++)
        Htr[a] += Hbig.a[(size_t)r * hx + a] * rbig[r];
  }
  // :397 measurement_compress_inplace — StateHelper.cpp:631-643
  if (Hc && rc && n_c) {
    if (Hbig.r > Hbig.c) {
      givens_AB(Hbig, rbig);
      *n_c = hx;
    } else {
      *n_c = Hbig.r;
    }
    for (int i = 0; i < *n_c; i++) {
      rc[i] = rbig[i];
      memcpy(Hc + (size_t)i * hx, &Hbig.a[(size_t)i * hx], sizeof(double) * hx);
    }
  }
  return total_res;
}

int orc_scan_linsys(const ml_scan_consts *c, const double *P, const float *map_xyzi, int M, const float *scan_L,
                    const float *scan_M, int N, const int32_t *knn_idx_in, double *HtH, double *Htr, ml_counters *cnt,
                    ml_point_row *rows_out, double *h_rows_out, double *Hc, double *rc, int *n_c, int threads) {
  return scan_linsys_core(c, P, map_xyzi, M, scan_L, scan_M, N, knn_idx_in, nullptr, nullptr, HtH, Htr, cnt, rows_out, h_rows_out, Hc,
                          rc, n_c, threads);
}
// same, with the neighbours given by coordinates (what KD_TREE::Nearest_Search returns): nb = N x k x float[4],
// nb_ok[i] = LidarHelper::get_neighbors' verdict. Used to time the reference tree + restated rows as one CPU path.
int orc_scan_linsys_nb(const ml_scan_consts *c, const double *P, const float *scan_L, int N, const float *nb, const uint8_t *nb_ok,
                       double *HtH, double *Htr, ml_counters *cnt, double *Hc, double *rc, int *n_c, int threads) {
  return scan_linsys_core(c, P, nullptr, 0, scan_L, nullptr, N, nullptr, nb, nb_ok, HtH, Htr, cnt, nullptr, nullptr, Hc, rc, n_c,
                          threads);
}

// ---------------------------------------------------------------------------------------------------------
// EKF update restricted to the marginal of H_order: StateHelper::EKFUpdate, StateHelper.cpp:121-200, with
// state->cov := P (n x n), H (m x n), R = I. dx = P H^T S^-1 res, Pnew = P - P H^T S^-1 H P.
// ---------------------------------------------------------------------------------------------------------
int orc_ekf_update(const double *P, int n, const double *H, const double *res, int m, double *dx, double *Pnew) {
  Mat Ma(n, m), S(m, m);
  for (int i = 0; i < n; i++)
    for (int j = 0; j < m; j++) {
      double s = 0;
      for (int l = 0; l < n; l++)
        s += P[(size_t)i * n + l] * H[(size_t)j * n + l];
      Ma(i, j) = s;
    }
  for (int i = 0; i < m; i++)
    for (int j = 0; j < m; j++) {
      double s = 0;
      for (int l = 0; l < n; l++)
        s += H[(size_t)i * n + l] * Ma(l, j);
      S(i, j) = s + (i == j ? 1.0 : 0.0);
    }
  if (!invert(S))
    return -1;
  Mat K(n, m);
  for (int i = 0; i < n; i++)
    for (int j = 0; j < m; j++) {
      double s = 0;
      for (int l = 0; l < m; l++)
        s += Ma(i, l) * S(l, j);
      K(i, j) = s;
    }
  for (int i = 0; i < n; i++) {
    double s = 0;
    for (int j = 0; j < m; j++)
      s += K(i, j) * res[j];
    dx[i] = s;
  }
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      double s = 0;
      for (int l = 0; l < m; l++)
        s += K(i, l) * Ma(j, l);
      Pnew[(size_t)i * n + j] = P[(size_t)i * n + j] - s;
    }
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// A0  LiDARData::LiDARData — LidarTypes.cpp:26-56. fp32 norm as Eigen evaluates Vector3f::norm():
// sqrt(x*x + (y*y + z*z)) (fixed-size redux unrolls as a binary tree). Returns the number kept.
// ---------------------------------------------------------------------------------------------------------
int orc_range_filter(const float *xyz, int stride, int n, float t_rel, double max_range, double min_range, float *out_xyzi) {
  int m = 0;
  for (int i = 0; i < n; i++) {
    const float *p = xyz + (size_t)i * stride;
    float yz = p[1] * p[1] + p[2] * p[2];
    float nn = std::sqrt(p[0] * p[0] + yz);
    if ((double)nn > max_range || (double)nn < min_range) // :39
      continue;
    out_xyzi[4 * (size_t)m + 0] = p[0];
    out_xyzi[4 * (size_t)m + 1] = p[1];
    out_xyzi[4 * (size_t)m + 2] = p[2];
    out_xyzi[4 * (size_t)m + 3] = t_rel; // :47
    m++;
  }
  return m;
}

// ---------------------------------------------------------------------------------------------------------
// A3  pcl::transformPointCloud(in, out, Matrix4f) — LidarHelper.cpp:192,203. PCL 1.10 (the noetic image,
// Dockerfile_noetic_20_04:1) evaluates, per output coordinate r: m[r][0]*x + (m[r][1]*y + (m[r][2]*z + m[r][3]))
// in fp32 (pcl/common/impl/transforms.hpp, SSE Transformer<float>::se3). UNPINNED: PCL is not in the tree.
// T row-major 4x4. intensity is copied.
// ---------------------------------------------------------------------------------------------------------
void orc_transform_f32(const float *in_xyzi, int n, const float *T, float *out_xyzi) {
  for (int i = 0; i < n; i++) {
    const float *p = in_xyzi + 4 * (size_t)i;
    float x = p[0], y = p[1], z = p[2];
    float o[3];
    for (int r = 0; r < 3; r++) {
      float a = T[4 * r + 2] * z + T[4 * r + 3];
      float b = T[4 * r + 1] * y + a;
      o[r] = T[4 * r + 0] * x + b;
    }
    float *q = out_xyzi + 4 * (size_t)i;
    q[0] = o[0], q[1] = o[1], q[2] = o[2], q[3] = p[3];
  }
}
// pcl::transformPointCloud(in, out, Matrix4d) — LidarHelper.cpp:261: generic Transformer<double>:
// float(((m0*x + m1*y) + m2*z) + m3) evaluated in fp64. UNPINNED.
void orc_transform_f64(const float *in_xyzi, int n, const double *T, float *out_xyzi) {
  for (int i = 0; i < n; i++) {
    const float *p = in_xyzi + 4 * (size_t)i;
    double x = p[0], y = p[1], z = p[2];
    float *q = out_xyzi + 4 * (size_t)i;
    float w = p[3];
    for (int r = 0; r < 3; r++)
      q[r] = (float)(((T[4 * r + 0] * x + T[4 * r + 1] * y) + T[4 * r + 2] * z) + T[4 * r + 3]);
    q[3] = w;
  }
}

// ---------------------------------------------------------------------------------------------------------
// A2  pcl::VoxelGrid<PointXYZI>::applyFilter with a cubic leaf (LidarHelper.cpp:89-101), PCL 1.10
// filters/impl/voxel_grid.hpp: bounds by getMinMax3D, inverse leaf in fp32, ijk = floor(p * inv_leaf) - min_b,
// idx = i + j*div_x + k*div_x*div_y, std::sort by idx, one centroid (all fields, fp32 sums / count) per run of
// equal idx, output in ascending idx. If dx*dy*dz overflows int32 PCL warns and copies the input.
// UNPINNED: PCL is not in the tree. Order of points inside a voxel follows std::sort on {idx, point index}
// compared by idx only, as PCL does (same libstdc++ introsort here).
// ---------------------------------------------------------------------------------------------------------
struct cloud_point_index_idx {
  unsigned int idx;
  unsigned int cloud_point_index;
  bool operator<(const cloud_point_index_idx &p) const { return idx < p.idx; }
};
int orc_voxel_downsample(const float *in_xyzi, int n, float leaf, float *out_xyzi, uint32_t *out_voxel_idx) {
  if (n == 0)
    return 0;
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = 0; i < n; i++)
    for (int a = 0; a < 3; a++) {
      mn[a] = std::min(mn[a], in_xyzi[4 * (size_t)i + a]);
      mx[a] = std::max(mx[a], in_xyzi[4 * (size_t)i + a]);
    }
  float inv = 1.0f / leaf;
  int64_t dx = (int64_t)((mx[0] - mn[0]) * inv) + 1, dy = (int64_t)((mx[1] - mn[1]) * inv) + 1,
          dz = (int64_t)((mx[2] - mn[2]) * inv) + 1;
  if (dx * dy * dz > (int64_t)std::numeric_limits<int32_t>::max()) {
    memcpy(out_xyzi, in_xyzi, sizeof(float) * 4 * (size_t)n);
    return n;
  }
  int minb[3], maxb[3], divb[3];
  for (int a = 0; a < 3; a++) {
    minb[a] = (int)std::floor(mn[a] * inv);
    maxb[a] = (int)std::floor(mx[a] * inv);
    divb[a] = maxb[a] - minb[a] + 1;
  }
  int mul[3] = {1, divb[0], divb[0] * divb[1]};
  std::vector<cloud_point_index_idx> iv(n);
  for (int i = 0; i < n; i++) {
    const float *p = in_xyzi + 4 * (size_t)i;
    int i0 = (int)std::floor(p[0] * inv) - minb[0];
    int i1 = (int)std::floor(p[1] * inv) - minb[1];
    int i2 = (int)std::floor(p[2] * inv) - minb[2];
    iv[i].idx = (unsigned)(i0 * mul[0] + i1 * mul[1] + i2 * mul[2]);
    iv[i].cloud_point_index = (unsigned)i;
  }
  std::sort(iv.begin(), iv.end());
  int m = 0;
  size_t i = 0;
  while (i < iv.size()) {
    size_t j = i;
    float s[4] = {0, 0, 0, 0};
    while (j < iv.size() && iv[j].idx == iv[i].idx) {
      const float *p = in_xyzi + 4 * (size_t)iv[j].cloud_point_index;
      for (int a = 0; a < 4; a++)
        s[a] += p[a];
      j++;
    }
    float cntf = (float)(j - i);
    for (int a = 0; a < 4; a++)
      out_xyzi[4 * (size_t)m + a] = s[a] / cntf;
    if (out_voxel_idx)
      out_voxel_idx[m] = iv[i].idx;
    m++;
    i = j;
  }
  return m;
}

// ---------------------------------------------------------------------------------------------------------
// A11 KD_TREE::Add_Points(points, downsample_on) — ikd_Tree.cpp:401-479 — restated on a flat point list so that
// map indices are defined: `map` holds the live points in order; new points are processed in input order; per
// point the box is [floor(x/ds)*ds, +ds) per axis in fp32 (:411-416), its centre min + (max-min)/2.0 (:417-419,
// evaluated in double then stored to float), candidates = live points inside the half-open box (Search_by_range,
// :1050-1053); the winner is the one closest to the centre, an existing point winning only if strictly closer
// (:422-430); if more than one candidate existed or the new point won, every other candidate of the box dies
// (:432-436): a winning new point is appended to the list, a winning existing point keeps its place (the tree has no
// list order; this is the flat-list convention that defines map indices). Without down-sampling points are appended.
// Output: the live list = surviving existing points in their old order, then surviving new points in input order.
// Returns the live count. cap = capacity of `out`.
// Ties among pre-existing points: the reference keeps whichever its tree traversal meets first (not a function of
// the point set); here the LAST in list order wins so the rule is uniform ("later wins").
// ---------------------------------------------------------------------------------------------------------
int orc_map_insert(const float *map_xyzi, int M, const float *add_xyzi, int n, int downsample, float ds, float *out, int cap) {
  struct Pt {
    float v[4];
    bool live;
  };
  std::vector<Pt> pts;
  pts.reserve((size_t)M + n);
  for (int i = 0; i < M; i++) {
    Pt p;
    memcpy(p.v, map_xyzi + 4 * (size_t)i, 16);
    p.live = true;
    pts.push_back(p);
  }
  if (!downsample) {
    for (int i = 0; i < n; i++) {
      Pt p;
      memcpy(p.v, add_xyzi + 4 * (size_t)i, 16);
      p.live = true;
      pts.push_back(p);
    }
  } else {
    // voxel key -> indices of live points with that key (key = floor(x/ds) per axis, the same fp32 expression)
    struct Key {
      int64_t a, b, c;
      bool operator<(const Key &o) const { return a != o.a ? a < o.a : (b != o.b ? b < o.b : c < o.c); }
    };
    auto keyof = [&](const float *p) {
      Key k;
      k.a = (int64_t)std::floor(p[0] / ds);
      k.b = (int64_t)std::floor(p[1] / ds);
      k.c = (int64_t)std::floor(p[2] / ds);
      return k;
    };
    std::map<Key, std::vector<int>> buckets;
    for (int i = 0; i < (int)pts.size(); i++)
      buckets[keyof(pts[i].v)].push_back(i);
    auto find_or_make = [&](const Key &k) -> std::vector<int> * { return &buckets[k]; };
    for (int i = 0; i < n; i++) {
      const float *np = add_xyzi + 4 * (size_t)i;
      float bmin[3], bmax[3], mid[3];
      for (int a = 0; a < 3; a++) {
        bmin[a] = std::floor(np[a] / ds) * ds;                    // :411
        bmax[a] = bmin[a] + ds;                                   // :412
        mid[a] = (float)(bmin[a] + (bmax[a] - bmin[a]) / 2.0);    // :417
      }
      std::vector<int> *bk = find_or_make(keyof(np));
      // Search_by_range: live points of this voxel inside the half-open box
      std::vector<int> storage;
      for (int id : *bk)
        if (pts[id].live) {
          const float *e = pts[id].v;
          bool in = true;
          for (int a = 0; a < 3; a++)
            in = in && (bmin[a] <= e[a] && bmax[a] > e[a]);
          if (in)
            storage.push_back(id);
        }
      float min_dist = calc_dist(np, mid); // :421
      int win = -1;                        // -1 = the new point
      for (int id : storage) {             // :423-429 (later wins ties among existing: see header)
        float td = calc_dist(pts[id].v, mid);
        if (td < min_dist || (win >= 0 && td == min_dist)) {
          min_dist = td;
          win = id;
        }
      }
      if (storage.size() > 1 || win < 0) { // :432
        for (int id : storage)
          if (id != win)
            pts[id].live = false; // Delete_by_range (a surviving existing point keeps its place in the list)
        if (win < 0) {
          Pt w;
          memcpy(w.v, np, 16);
          w.live = true;
          pts.push_back(w); // Add_by_point
          bk->push_back((int)pts.size() - 1);
        }
      }
    }
  }
  int m = 0;
  for (auto &p : pts)
    if (p.live) {
      if (m < cap)
        memcpy(out + 4 * (size_t)m, p.v, 16);
      m++;
    }
  return m;
}

// ---------------------------------------------------------------------------------------------------------
// A12 LidarHelper::propagate_map_to_newest_clone — LidarHelper.cpp:209-266 (pose look-ups excluded: T is an input).
// flatten -> [VoxelGrid centroids] (:214) -> decay filters (:219-226; fp32 norm as Eigen: sqrt(x*x+(y*y+z*z)))
// -> transformPointCloud with a Matrix4d (:261) -> Build.
// ---------------------------------------------------------------------------------------------------------
int orc_reanchor(const float *map_xyzi, int M, const double *T, double now_minus_FT, double decay_time, double decay_dist,
                 int do_ds, float ds, float *out) {
  std::vector<float> a(4 * (size_t)std::max(M, 1)), b(4 * (size_t)std::max(M, 1));
  int m = M;
  if (do_ds)
    m = orc_voxel_downsample(map_xyzi, M, ds, a.data(), nullptr);
  else
    memcpy(a.data(), map_xyzi, sizeof(float) * 4 * (size_t)M);
  int kept = 0;
  for (int i = 0; i < m; i++) {
    const float *p = &a[4 * (size_t)i];
    if ((double)p[3] + decay_time < now_minus_FT) // pt.intensity + FT + map_decay_time < state->time (:220)
      continue;
    float yz = p[1] * p[1] + p[2] * p[2];
    float nn = std::sqrt(p[0] * p[0] + yz);
    if ((double)nn > decay_dist) // :222
      continue;
    memcpy(&b[4 * (size_t)kept], p, 16);
    kept++;
  }
  orc_transform_f64(b.data(), kept, T, out);
  return kept;
}

// ---------------------------------------------------------------------------------------------------------
// A1  LidarHelper::remove_motion_blur loop body — LidarHelper.cpp:71-85 — given the per-point transforms:
// pcl::transformPoint(p, Affine3f T) = T.linear()*p + T.translation() evaluated by Eigen as
// ((r0*x + r1*y) + r2*z) + t in fp32. T_LitoL: n x 12 (row-major 3x4). UNPINNED (no reachable reference path:
// the guard at :54-58 is inverted so the loop never runs).
// ---------------------------------------------------------------------------------------------------------
void orc_deskew(float *xyzi, int n, const float *T_LitoL) {
  for (int i = 0; i < n; i++) {
    float *p = xyzi + 4 * (size_t)i;
    const float *T = T_LitoL + 12 * (size_t)i;
    float x = p[0], y = p[1], z = p[2];
    for (int r = 0; r < 3; r++)
      p[r] = ((T[4 * r] * x + T[4 * r + 1] * y) + T[4 * r + 2] * z) + T[4 * r + 3];
  }
}

int orc_sizeof_consts() { return (int)sizeof(ml_scan_consts); }
int orc_sizeof_row() { return (int)sizeof(ml_point_row); }
int orc_sizeof_counters() { return (int)sizeof(ml_counters); }
}
