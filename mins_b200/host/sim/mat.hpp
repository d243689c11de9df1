// Small dense algebra for the closed-loop harness (host C++14, fp64, row-major). Not part of the LiDAR hot path: it carries the
// minimal EKF core that stands in for mins/src/state/{State,StateHelper,Propagator}.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace mins_sim {

struct Mat {
  int r = 0, c = 0;
  std::vector<double> a;
  Mat() {}
  Mat(int r_, int c_) : r(r_), c(c_), a((size_t)r_ * c_, 0.0) {}
  double &operator()(int i, int j) { return a[(size_t)i * c + j]; }
  double operator()(int i, int j) const { return a[(size_t)i * c + j]; }
  static Mat eye(int n) {
    Mat m(n, n);
    for (int i = 0; i < n; i++)
      m(i, i) = 1.0;
    return m;
  }
};

inline Mat mul(const Mat &A, const Mat &B) {
  Mat C(A.r, B.c);
  for (int i = 0; i < A.r; i++)
    for (int k = 0; k < A.c; k++) {
      const double x = A(i, k);
      if (x == 0.0)
        continue;
      const double *b = &B.a[(size_t)k * B.c];
      double *cr = &C.a[(size_t)i * C.c];
      for (int j = 0; j < B.c; j++)
        cr[j] += x * b[j];
    }
  return C;
}
inline Mat transpose(const Mat &A) {
  Mat T(A.c, A.r);
  for (int i = 0; i < A.r; i++)
    for (int j = 0; j < A.c; j++)
      T(j, i) = A(i, j);
  return T;
}
// general inverse by Gauss-Jordan with partial pivoting (small matrices: 3x3 .. 9x9)
inline Mat inverse(const Mat &Ain) {
  const int n = Ain.r;
  Mat A = Ain, I = Mat::eye(n);
  for (int k = 0; k < n; k++) {
    int p = k;
    for (int i = k + 1; i < n; i++)
      if (std::fabs(A(i, k)) > std::fabs(A(p, k)))
        p = i;
    if (A(p, k) == 0.0)
      throw std::runtime_error("singular matrix");
    if (p != k)
      for (int j = 0; j < n; j++) {
        std::swap(A(k, j), A(p, j));
        std::swap(I(k, j), I(p, j));
      }
    const double d = 1.0 / A(k, k);
    for (int j = 0; j < n; j++) {
      A(k, j) *= d;
      I(k, j) *= d;
    }
    for (int i = 0; i < n; i++) {
      if (i == k)
        continue;
      const double f = A(i, k);
      if (f == 0.0)
        continue;
      for (int j = 0; j < n; j++) {
        A(i, j) -= f * A(k, j);
        I(i, j) -= f * I(k, j);
      }
    }
  }
  return I;
}
// X = S^-1 B for symmetric positive definite S (Cholesky); returns false if S is not positive definite
inline bool chol_solve(const Mat &S, const Mat &B, Mat &X) {
  const int n = S.r;
  Mat L(n, n);
  for (int j = 0; j < n; j++) {
    double d = S(j, j);
    for (int k = 0; k < j; k++)
      d -= L(j, k) * L(j, k);
    if (!(d > 0.0))
      return false;
    L(j, j) = std::sqrt(d);
    for (int i = j + 1; i < n; i++) {
      double s = S(i, j);
      for (int k = 0; k < j; k++)
        s -= L(i, k) * L(j, k);
      L(i, j) = s / L(j, j);
    }
  }
  X = B;
  for (int col = 0; col < B.c; col++) {
    for (int i = 0; i < n; i++) {
      double s = X(i, col);
      for (int k = 0; k < i; k++)
        s -= L(i, k) * X(k, col);
      X(i, col) = s / L(i, i);
    }
    for (int i = n - 1; i >= 0; i--) {
      double s = X(i, col);
      for (int k = i + 1; k < n; k++)
        s -= L(k, i) * X(k, col);
      X(i, col) = s / L(i, i);
    }
  }
  return true;
}

// ---- 3-vectors and 3x3 matrices as plain arrays (row-major) ----------------------------------------------------------
inline void m3_mul(const double *A, const double *B, double *C) {
  double t[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      t[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
  std::memcpy(C, t, sizeof(t));
}
inline void m3_T(const double *A, double *T) {
  double t[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      t[3 * i + j] = A[3 * j + i];
  std::memcpy(T, t, sizeof(t));
}
inline void m3_vec(const double *A, const double *v, double *o) {
  double t[3];
  for (int i = 0; i < 3; i++)
    t[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
  std::memcpy(o, t, sizeof(t));
}
inline void m3_Tvec(const double *A, const double *v, double *o) {
  double t[3];
  for (int i = 0; i < 3; i++)
    t[i] = A[i] * v[0] + A[3 + i] * v[1] + A[6 + i] * v[2];
  std::memcpy(o, t, sizeof(t));
}
inline void m3_eye(double *A) {
  for (int i = 0; i < 9; i++)
    A[i] = (i % 4 == 0) ? 1.0 : 0.0;
}
inline void skew(const double *w, double *S) {
  S[0] = 0, S[1] = -w[2], S[2] = w[1];
  S[3] = w[2], S[4] = 0, S[5] = -w[0];
  S[6] = -w[1], S[7] = w[0], S[8] = 0;
}
inline double norm3(const double *v) { return std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }
inline double det3(const double *A) {
  return A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) + A[2] * (A[3] * A[7] - A[4] * A[6]);
}
inline bool m3_inv(const double *A, double *I) {
  const double d = det3(A);
  if (d == 0.0)
    return false;
  const double id = 1.0 / d;
  double t[9];
  t[0] = (A[4] * A[8] - A[5] * A[7]) * id, t[1] = (A[2] * A[7] - A[1] * A[8]) * id, t[2] = (A[1] * A[5] - A[2] * A[4]) * id;
  t[3] = (A[5] * A[6] - A[3] * A[8]) * id, t[4] = (A[0] * A[8] - A[2] * A[6]) * id, t[5] = (A[2] * A[3] - A[0] * A[5]) * id;
  t[6] = (A[3] * A[7] - A[4] * A[6]) * id, t[7] = (A[1] * A[6] - A[0] * A[7]) * id, t[8] = (A[0] * A[4] - A[1] * A[3]) * id;
  std::memcpy(I, t, sizeof(t));
  return true;
}

// Rotation R maximising trace(R^T W) (Kabsch / Umeyama without scale: ov_eval AlignTrajectory se3), via the symmetric
// eigen-decomposition of W^T W by cyclic Jacobi: W = U S V^T, R = U diag(1, 1, det(U V^T)) V^T.
inline void kabsch(const double *W, double *R) {
  double A[9], V[9];
  double Wt[9];
  m3_T(W, Wt);
  m3_mul(Wt, W, A);
  m3_eye(V);
  for (int sweep = 0; sweep < 60; sweep++) {
    double off = std::fabs(A[1]) + std::fabs(A[2]) + std::fabs(A[5]);
    if (off < 1e-300)
      break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++) {
        const double apq = A[3 * p + q];
        if (apq == 0.0)
          continue;
        const double th = (A[3 * q + q] - A[3 * p + p]) / (2.0 * apq);
        const double t = (th >= 0 ? 1.0 : -1.0) / (std::fabs(th) + std::sqrt(th * th + 1.0));
        const double cs = 1.0 / std::sqrt(t * t + 1.0), sn = t * cs;
        double G[9];
        m3_eye(G);
        G[3 * p + p] = cs, G[3 * q + q] = cs, G[3 * p + q] = sn, G[3 * q + p] = -sn;
        double Gt[9], tmp[9];
        m3_T(G, Gt);
        m3_mul(Gt, A, tmp);
        m3_mul(tmp, G, A);
        m3_mul(V, G, V);
      }
  }
  // singular values and U = W V S^-1 (columns); a vanishing singular value gets the cross product of the others
  double U[9], WV[9];
  m3_mul(W, V, WV);
  double s[3];
  for (int j = 0; j < 3; j++)
    s[j] = std::sqrt(std::max(0.0, A[3 * j + j]));
  int order[3] = {0, 1, 2};
  std::sort(order, order + 3, [&](int x, int y) { return s[x] > s[y]; });
  for (int jj = 0; jj < 2; jj++) {
    const int j = order[jj];
    const double n = std::sqrt(WV[j] * WV[j] + WV[3 + j] * WV[3 + j] + WV[6 + j] * WV[6 + j]);
    for (int i = 0; i < 3; i++)
      U[3 * i + j] = n > 0 ? WV[3 * i + j] / n : (i == j ? 1.0 : 0.0);
  }
  {
    const int a = order[0], b = order[1], cc = order[2];
    const double ua[3] = {U[a], U[3 + a], U[6 + a]}, ub[3] = {U[b], U[3 + b], U[6 + b]};
    double uc[3] = {ua[1] * ub[2] - ua[2] * ub[1], ua[2] * ub[0] - ua[0] * ub[2], ua[0] * ub[1] - ua[1] * ub[0]};
    // orientation: make det(U) det(V) consistent -> R = U D V^T with D fixing the reflection on the smallest singular value
    double Vc[3] = {V[cc], V[3 + cc], V[6 + cc]}, Va[3] = {V[a], V[3 + a], V[6 + a]}, Vb[3] = {V[b], V[3 + b], V[6 + b]};
    double vx[3] = {Va[1] * Vb[2] - Va[2] * Vb[1], Va[2] * Vb[0] - Va[0] * Vb[2], Va[0] * Vb[1] - Va[1] * Vb[0]};
    const double sgn = (vx[0] * Vc[0] + vx[1] * Vc[1] + vx[2] * Vc[2]) >= 0 ? 1.0 : -1.0; // handedness of (Va, Vb, Vc)
    for (int i = 0; i < 3; i++)
      U[3 * i + cc] = sgn * uc[i]; // (ua, ub, U_c) gets the same handedness: det(U V^T) = +1
  }
  double Vt[9];
  m3_T(V, Vt);
  m3_mul(U, Vt, R);
}

} // namespace mins_sim
