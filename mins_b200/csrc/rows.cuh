// Per-point measurement rows and their fixed-order reduction.
// Reference: mins/src/update/lidar/UpdaterLidar.cpp:279-368 (point loop), LidarHelper.cpp:291-328 (plane),
// mins/src/state/StateHelper.cpp:645-680 (nullspace projection), mins/src/update/UpdaterStatistics.cpp:94-117 (chi2).
//
// What the reference does per point and how it is evaluated here (derivations in DESIGN.md §"closed forms"):
//  * H (6 x total_hx) has a single non-zero row h^T (UpdaterLidar.cpp:342-350) and h = T^T g with the per-scan
//    constant T (18 x total_hx, built from dTdxI/dTdxA/calib columns) and the per-point 18-vector
//    g = [n dpfinM_dI | n dpfinM_dA | n dpfinM_dcalib]. So h^T P h = g^T (T P T^T) g and
//    sum_i w_i h_i h_i^T = T^T (sum_i w_i g_i g_i^T) T: all per-point work is 18-wide instead of total_hx-wide.
//  * The Givens nullspace projection of the 6x3 H_f (StateHelper.cpp:645-680) leaves H' = v h^T (rank one) with
//    ||v||^2 = 1/(1+t), v^T r' = e/(1+t), ||r'||^2 = RSS_nb + e^2/(1+t), where t = a5^T (A_w^T A_w)^-1 a5 is the
//    leverage of the scan-point row against the k whitened neighbour rows, e its deleted residual and RSS_nb the
//    residual sum of squares of the neighbour rows alone (recursive-least-squares identities).
//  * chi2 = r'^T (H' P H'^T + I)^-1 r' = ||r'||^2 - s (v^T r')^2 / (1 + s ||v||^2), s = h^T P h (Sherman-Morrison).
//  * plane: Householder QR of the k x 3 neighbour matrix (same least-squares solution as colPivHouseholderQr for
//    the full-rank matrices that pass the condition gate); cond(A) from the eigenvalues of A^T A.
#pragma once
#include "common.cuh"

namespace ml {

__device__ __forceinline__ void d_mat3_vec(const double *A, const double *v, double *o) {
#pragma unroll
  for (int i = 0; i < 3; i++)
    o[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
}
__device__ __forceinline__ void d_mat3T_vec(const double *A, const double *v, double *o) {
#pragma unroll
  for (int i = 0; i < 3; i++)
    o[i] = A[i] * v[0] + A[3 + i] * v[1] + A[6 + i] * v[2];
}
__device__ __forceinline__ void d_cross(const double *a, const double *b, double *o) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}

// eigenvalues (min, max) of a symmetric 3x3 matrix [a00 a01 a02; . a11 a12; . . a22], trigonometric closed form
__device__ __forceinline__ void sym3_eig_minmax(double a00, double a01, double a02, double a11, double a12, double a22, double *lmin,
                                                double *lmax) {
  const double p1 = a01 * a01 + a02 * a02 + a12 * a12;
  const double q = (a00 + a11 + a22) / 3.0;
  const double b00 = a00 - q, b11 = a11 - q, b22 = a22 - q;
  const double p2 = b00 * b00 + b11 * b11 + b22 * b22 + 2.0 * p1;
  if (p2 <= 0.0) {
    *lmin = *lmax = q;
    return;
  }
  const double p = sqrt(p2 / 6.0);
  const double ip = 1.0 / p;
  const double c00 = b00 * ip, c11 = b11 * ip, c22 = b22 * ip, c01 = a01 * ip, c02 = a02 * ip, c12 = a12 * ip;
  double r = 0.5 * (c00 * (c11 * c22 - c12 * c12) - c01 * (c01 * c22 - c12 * c02) + c02 * (c01 * c12 - c11 * c02));
  r = fmin(1.0, fmax(-1.0, r));
  const double phi = acos(r) / 3.0;
  *lmax = q + 2.0 * p * cos(phi);
  *lmin = q + 2.0 * p * cos(phi + 2.0943951023931953);
}

// T (18 x hx) and Q = T P T^T from the scan constants. One block.
__global__ void __launch_bounds__(256) k_prep_consts(DevConsts *__restrict__ dc, const double *__restrict__ P) {
  __shared__ double TP[ML_G * ML_MAX_HX];
  const ml_scan_consts &c = dc->c;
  const int hx = c.total_hx;
  for (int e = threadIdx.x; e < ML_G * ML_MAX_HX; e += blockDim.x)
    dc->T[e] = 0.0;
  __syncthreads();
  // one thread per (block, row, col) would race on overlapping columns (I and A clones may coincide): serialise per row
  if (threadIdx.x < ML_G) {
    const int r = threadIdx.x;
    double *Tr = dc->T + r * ML_MAX_HX;
    if (r < 6) {
      for (int b = 0; b < c.nI; b++)
        for (int cc = 0; cc < c.dTdxI[b].width; cc++)
          Tr[c.dTdxI[b].col + cc] += c.dTdxI[b].D[6 * r + cc];
    } else if (r < 12) {
      for (int b = 0; b < c.nA; b++)
        for (int cc = 0; cc < c.dTdxA[b].width; cc++)
          Tr[c.dTdxA[b].col + cc] += c.dTdxA[b].D[6 * (r - 6) + cc];
    } else if (c.calib_col >= 0) {
      Tr[c.calib_col + (r - 12)] += 1.0;
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < ML_G * hx; e += blockDim.x) {
    const int r = e / hx, j = e - r * hx;
    double s = 0;
    for (int l = 0; l < hx; l++)
      s += dc->T[r * ML_MAX_HX + l] * P[(size_t)l * hx + j];
    TP[r * ML_MAX_HX + j] = s;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < ML_G * ML_G; e += blockDim.x) {
    const int r = e / ML_G, q = e - r * ML_G;
    double s = 0;
    for (int l = 0; l < hx; l++)
      s += TP[r * ML_MAX_HX + l] * dc->T[q * ML_MAX_HX + l];
    dc->Q[e] = s;
  }
}

constexpr int ROWS_BLOCK = 128;

template <int K>
__global__ void __launch_bounds__(ROWS_BLOCK)
    k_rows(const ScanState *__restrict__ st, const float4 *__restrict__ scan_L, const float4 *__restrict__ map_pts,
           const int32_t *__restrict__ knn_idx, const uint8_t *__restrict__ knn_ok, const DevConsts *__restrict__ dc,
           PointRow *__restrict__ rows) {
  __shared__ double sQ[ML_G * ML_G];
  __shared__ ml_scan_consts sc_s;
  {
    const int nw = (int)(sizeof(ml_scan_consts) / sizeof(int));
    const int *src = reinterpret_cast<const int *>(&dc->c);
    int *dst = reinterpret_cast<int *>(&sc_s);
    for (int e = threadIdx.x; e < nw; e += blockDim.x)
      dst[e] = src[e];
    for (int e = threadIdx.x; e < ML_G * ML_G; e += blockDim.x)
      sQ[e] = dc->Q[e];
  }
  __syncthreads();
  const ml_scan_consts &c = sc_s;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= st->n_scan)
    return;
  PointRow out;
#pragma unroll
  for (int a = 0; a < ML_G; a++)
    out.g[a] = 0.0;
  out.w = out.vr = out.chi2 = out.res = out.raw_n = out.rss = 0.0;
  out.plane[0] = out.plane[1] = out.plane[2] = out.plane[3] = 0.0;
  out.pad = 0;
  out.status = 1;
  if (!knn_ok[i]) {
    rows[i] = out;
    return;
  }
  // ---- neighbours, A = k x 3, d = mean range (LidarHelper.cpp:293-302) ----
  double A[K][3];
  double dmean = 0.0;
#pragma unroll
  for (int j = 0; j < K; j++) {
    const float4 p = __ldg(&map_pts[knn_idx[(size_t)i * K + j]]);
    A[j][0] = (double)p.x, A[j][1] = (double)p.y, A[j][2] = (double)p.z;
    dmean += sqrt(A[j][0] * A[j][0] + A[j][1] * A[j][1] + A[j][2] * A[j][2]);
  }
  dmean /= (double)K;
  // ---- condition number gate (LidarHelper.cpp:305-312): cond(A)^2 = lmax/lmin of A^T A ----
  {
    double g00 = 0, g01 = 0, g02 = 0, g11 = 0, g12 = 0, g22 = 0;
#pragma unroll
    for (int j = 0; j < K; j++) {
      g00 += A[j][0] * A[j][0], g01 += A[j][0] * A[j][1], g02 += A[j][0] * A[j][2];
      g11 += A[j][1] * A[j][1], g12 += A[j][1] * A[j][2], g22 += A[j][2] * A[j][2];
    }
    double lmin, lmax;
    sym3_eig_minmax(g00, g01, g02, g11, g12, g22, &lmin, &lmax);
    // condA > max  <=>  lmax > max^2 * lmin  (lmin <= 0 after rounding means singular: reject, as inf/NaN > max
    // would... NaN compares false in the reference, but a rank-deficient 5x3 A never passes the later gates)
    if (!(lmin > 0.0) || lmax > c.plane_max_condi * c.plane_max_condi * lmin) {
      out.status = 2;
      rows[i] = out;
      return;
    }
  }
  // keep the raw neighbour coordinates for the residuals
  double PN[K][3];
#pragma unroll
  for (int j = 0; j < K; j++) {
    PN[j][0] = A[j][0], PN[j][1] = A[j][1], PN[j][2] = A[j][2];
  }
  // ---- Householder QR of A, applied to b = -d*1 (LidarHelper.cpp:302,315) ----
  double b[K];
#pragma unroll
  for (int j = 0; j < K; j++)
    b[j] = -dmean;
  double Rd[3], vdiag[3], vn2[3];
#pragma unroll
  for (int cc = 0; cc < 3; cc++) {
    double sigma = 0.0;
#pragma unroll
    for (int r = cc; r < K; r++)
      sigma += A[r][cc] * A[r][cc];
    const double nrm = sqrt(sigma);
    const double alpha = A[cc][cc] > 0.0 ? -nrm : nrm;
    const double v0 = A[cc][cc] - alpha;
    const double vnorm2 = sigma - A[cc][cc] * A[cc][cc] + v0 * v0;
    Rd[cc] = alpha;
    vdiag[cc] = v0;
    vn2[cc] = vnorm2;
    const double inv2 = vnorm2 > 0.0 ? 2.0 / vnorm2 : 0.0;
#pragma unroll
    for (int jj = cc + 1; jj < 3; jj++) {
      double dot = v0 * A[cc][jj];
#pragma unroll
      for (int r = cc + 1; r < K; r++)
        dot += A[r][cc] * A[r][jj];
      const double f = dot * inv2;
      A[cc][jj] -= f * v0;
#pragma unroll
      for (int r = cc + 1; r < K; r++)
        A[r][jj] -= f * A[r][cc];
    }
    {
      double dot = v0 * b[cc];
#pragma unroll
      for (int r = cc + 1; r < K; r++)
        dot += A[r][cc] * b[r];
      const double f = dot * inv2;
      b[cc] -= f * v0;
#pragma unroll
      for (int r = cc + 1; r < K; r++)
        b[r] -= f * A[r][cc];
    }
  }
  // R = [Rd0 A01 A02; 0 Rd1 A12; 0 0 Rd2]; abc = R^-1 (Q^T b)[0:3]
  double n[3];
  n[2] = b[2] / Rd[2];
  n[1] = (b[1] - A[1][2] * n[2]) / Rd[1];
  n[0] = (b[0] - A[0][1] * n[1] - A[0][2] * n[2]) / Rd[0];
  out.plane[0] = n[0], out.plane[1] = n[1], out.plane[2] = n[2], out.plane[3] = dmean;
  const double nn = sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
  // ---- sanity check (LidarHelper.cpp:319-325) ----
  {
    const double pl0 = n[0] / nn, pl1 = n[1] / nn, pl2 = n[2] / nn, pl3 = dmean / nn;
    bool bad = false;
#pragma unroll
    for (int j = 0; j < K; j++)
      bad = bad || (fabs(pl0 * PN[j][0] + pl1 * PN[j][1] + pl2 * PN[j][2] + pl3) > c.plane_max_p2pd);
    if (bad || !(nn == nn)) {
      out.status = 2;
      rows[i] = out;
      return;
    }
  }
  // ---- scan point in the map frame, est and fej (UpdaterLidar.cpp:281-283) ----
  const float4 pl4 = scan_L[i];
  const double pfinL[3] = {(double)pl4.x, (double)pl4.y, (double)pl4.z};
  double t3[3], pfinM_est[3], pfinM_fej[3];
  d_mat3_vec(c.RLtoM_est, pfinL, t3);
#pragma unroll
  for (int a = 0; a < 3; a++)
    pfinM_est[a] = c.pLinM_est[a] + t3[a];
  d_mat3_vec(c.RLtoM_fej, pfinL, t3);
#pragma unroll
  for (int a = 0; a < 3; a++)
    pfinM_fej[a] = c.pLinM_fej[a] + t3[a];
  // ---- bad match gate + residual of the scan point (UpdaterLidar.cpp:296,313) ----
  const double res5 = -(n[0] * pfinM_est[0] + n[1] * pfinM_est[1] + n[2] * pfinM_est[2]) - dmean;
  if (res5 > c.plane_max_p2pd * 3.0) {
    out.status = 3;
    rows[i] = out;
    return;
  }
  // ---- g = n * [dpfinM_dI | dpfinM_dA | dpfinM_dcalib] (UpdaterLidar.cpp:320-329) ----
  double g[ML_G];
  {
    double nG[3], nI[3], w[3], v[3], cr[3];
    d_mat3T_vec(c.RGtoM_fej, n, nG); // (n^T RGtoM)^T
    d_mat3T_vec(c.RItoG_fej, nG, nI); // (n^T RGtoM RItoG)^T
    // -n^T RGtoM RItoG skew(w) = (w x nI)^T, w = RLtoI (pfinL - pIinL)
#pragma unroll
    for (int a = 0; a < 3; a++)
      v[a] = pfinL[a] - c.pIinL_fej[a];
    d_mat3_vec(c.RLtoI_fej, v, w);
    d_cross(w, nI, cr);
    g[0] = cr[0], g[1] = cr[1], g[2] = cr[2];
    g[3] = nG[0], g[4] = nG[1], g[5] = nG[2];
    // n^T RItoL skew(u) = (a2 x u)^T, a2 = RItoL^T n, u = RGtoA (pIinG + RItoG pLinI - pAinG + RLtoG pfinL)
    double a2[3], u[3], t1[3], t2[3];
    d_mat3T_vec(c.RItoL_fej, n, a2);
    d_mat3_vec(c.RItoG_fej, c.pLinI_fej, t1);
    d_mat3_vec(c.RLtoG_fej, pfinL, t2);
#pragma unroll
    for (int a = 0; a < 3; a++)
      v[a] = c.pIinG_fej[a] + t1[a] - c.pAinG_fej[a] + t2[a];
    d_mat3_vec(c.RGtoA_fej, v, u);
    d_cross(a2, u, cr);
    g[6] = cr[0], g[7] = cr[1], g[8] = cr[2];
    g[9] = -nG[0], g[10] = -nG[1], g[11] = -nG[2];
    // n^T (skew(c1) - skew(c2) + RLtoM skew(c3)) = (n x c1) - (n x c2) + (a3 x c3)
    double c1[3], c2[3], c3[3], a3[3], x1[3], x2[3], x3[3];
#pragma unroll
    for (int a = 0; a < 3; a++)
      v[a] = c.pIinG_fej[a] - c.pAinG_fej[a];
    d_mat3_vec(c.RGtoM_fej, v, c1);
#pragma unroll
    for (int a = 0; a < 3; a++)
      c3[a] = c.pIinL_fej[a] - pfinL[a];
    d_mat3_vec(c.RLtoM_fej, c3, c2);
    d_mat3T_vec(c.RLtoM_fej, n, a3);
    d_cross(n, c1, x1);
    d_cross(n, c2, x2);
    d_cross(a3, c3, x3);
#pragma unroll
    for (int a = 0; a < 3; a++) {
      g[12 + a] = x1[a] - x2[a] + x3[a];
      g[15 + a] = n[a] - a3[a];
    }
    if (c.calib_col < 0) {
#pragma unroll
      for (int a = 12; a < 18; a++)
        g[a] = 0.0;
    }
  }
  // ---- interpolation-error inflation (UpdaterLidar.cpp:332-337) and whitening scales (:353-354) ----
  double intr_std = 0.0;
  if (!c.at_clone_m || !c.at_clone_a)
    intr_std = sqrt(c.ie_o * (g[0] * g[0] + g[1] * g[1] + g[2] * g[2]) + c.ie_p * (g[3] * g[3] + g[4] * g[4] + g[5] * g[5]));
  const double raw_n = nn * c.raw_noise + intr_std;
  const double map_n = nn * c.map_noise + intr_std;
  // ---- neighbour residuals (UpdaterLidar.cpp:305-310), pushed through the same reflectors: Q^T res_nb ----
  double rn[K];
#pragma unroll
  for (int j = 0; j < K; j++)
    rn[j] = -(n[0] * PN[j][0] + n[1] * PN[j][1] + n[2] * PN[j][2]) - dmean;
#pragma unroll
  for (int cc = 0; cc < 3; cc++) {
    const double inv2 = vn2[cc] > 0.0 ? 2.0 / vn2[cc] : 0.0;
    double dot = vdiag[cc] * rn[cc];
#pragma unroll
    for (int r = cc + 1; r < K; r++)
      dot += A[r][cc] * rn[r];
    const double f = dot * inv2;
    rn[cc] -= f * vdiag[cc];
#pragma unroll
    for (int r = cc + 1; r < K; r++)
      rn[r] -= f * A[r][cc];
  }
  // beta_nb = R^-1 (Q^T res_nb)[0:3] (zero up to rounding: res_nb is the least-squares residual of the plane fit)
  double beta[3];
  beta[2] = rn[2] / Rd[2];
  beta[1] = (rn[1] - A[1][2] * beta[2]) / Rd[1];
  beta[0] = (rn[0] - A[0][1] * beta[1] - A[0][2] * beta[2]) / Rd[0];
  double rss_nb = 0.0;
#pragma unroll
  for (int r = 3; r < K; r++)
    rss_nb += rn[r] * rn[r];
  rss_nb /= (map_n * map_n);
  // leverage of the scan-point row: t = (map_n/raw_n)^2 || R^-T pfinM_fej ||^2
  double z[3];
  z[0] = pfinM_fej[0] / Rd[0];
  z[1] = (pfinM_fej[1] - A[0][1] * z[0]) / Rd[1];
  z[2] = (pfinM_fej[2] - A[0][2] * z[0] - A[1][2] * z[1]) / Rd[2];
  const double ratio = map_n / raw_n;
  const double tlev = ratio * ratio * (z[0] * z[0] + z[1] * z[1] + z[2] * z[2]);
  // deleted residual of the scan-point row (whitened)
  const double e5 = (res5 - (pfinM_fej[0] * beta[0] + pfinM_fej[1] * beta[1] + pfinM_fej[2] * beta[2])) / raw_n;
  const double wv = 1.0 / (1.0 + tlev);
  const double vr = e5 * wv;
  const double rss = rss_nb + e5 * e5 * wv;
  // ---- s = h^T P h = g_w^T Q g_w ----
  const double ir = 1.0 / raw_n;
#pragma unroll
  for (int a = 0; a < ML_G; a++)
    g[a] *= ir;
  double s = 0.0;
#pragma unroll
  for (int a = 0; a < ML_G; a++) {
    double acc = 0.0;
#pragma unroll
    for (int bq = 0; bq < ML_G; bq++)
      acc += sQ[a * ML_G + bq] * g[bq];
    s += g[a] * acc;
  }
  const double chi2 = rss - s * vr * vr / (1.0 + s * wv);
#pragma unroll
  for (int a = 0; a < ML_G; a++)
    out.g[a] = g[a];
  out.w = wv;
  out.vr = vr;
  out.chi2 = chi2;
  out.res = res5 / raw_n;
  out.raw_n = raw_n;
  out.rss = rss;
  out.status = (chi2 < c.chi2_thr) ? 0 : 4; // NaN -> reject (UpdaterStatistics.cpp:117)
  rows[i] = out;
}

// ------------------------------------------------------------------------------------------------
// Reduction: partial[b] = sum over a contiguous chunk of points, in index order (bit-reproducible).
// thread e < 171: G entry (i<=j); 171 <= e < 189: b entry; warp 6 (threads 192..223): counters.
// ------------------------------------------------------------------------------------------------
constexpr int RED_BLOCK = 224;
constexpr int RED_TILE = 64;

__global__ void __launch_bounds__(RED_BLOCK) k_reduce(const ScanState *__restrict__ st, const PointRow *__restrict__ rows, int nparts,
                                                     Partial *__restrict__ parts) {
  __shared__ double sg[RED_TILE][ML_G + 1];
  __shared__ double swg[RED_TILE][ML_G + 1];
  __shared__ double svr[RED_TILE];
  __shared__ double schi[RED_TILE], srss[RED_TILE];
  __shared__ int sstat[RED_TILE];
  const int n = st->n_scan;
  const int chunk = (n + nparts - 1) / nparts;
  const int p0 = blockIdx.x * chunk, p1 = min(n, p0 + chunk);
  const int e = threadIdx.x;
  int gi = 0, gj = 0;
  if (e < ML_GSYM) {
    int r = 0, rem = e;
    while (rem >= ML_G - r) {
      rem -= ML_G - r;
      r++;
    }
    gi = r;
    gj = r + rem;
  }
  double acc = 0.0;
  // counters (warp 6)
  int c_knn = 0, c_plane = 0, c_bad = 0, c_chi = 0, c_acc = 0;
  double s_chi = 0, s_chi2 = 0, s_res = 0, s_res2 = 0;
  for (int base = p0; base < p1; base += RED_TILE) {
    const int cnt = min(RED_TILE, p1 - base);
    for (int idx = threadIdx.x; idx < RED_TILE * (ML_G + 1); idx += RED_BLOCK) {
      const int p = idx / (ML_G + 1), a = idx - p * (ML_G + 1);
      if (p < cnt) {
        const PointRow *r = &rows[base + p];
        if (a < ML_G) {
          const double gv = r->g[a];
          const bool ok = r->status == 0;
          sg[p][a] = ok ? gv : 0.0;
          swg[p][a] = ok ? gv * r->w : 0.0;
        } else {
          svr[p] = r->status == 0 ? r->vr : 0.0;
          schi[p] = r->chi2;
          srss[p] = r->rss;
          sstat[p] = r->status;
        }
      }
    }
    __syncthreads();
    if (e < ML_GSYM) {
      for (int p = 0; p < cnt; p++)
        acc += swg[p][gi] * sg[p][gj];
    } else if (e < ML_GSYM + ML_G) {
      const int a = e - ML_GSYM;
      for (int p = 0; p < cnt; p++)
        acc += sg[p][a] * svr[p];
    } else if (e >= 192) {
      const int lane = e - 192;
      for (int h = 0; h < RED_TILE; h += 32) {
        const int p = h + lane;
        const int stt = p < cnt ? sstat[p] : -1;
        const bool tested = (stt == 0 || stt == 4) && (schi[p < cnt ? p : 0] == schi[p < cnt ? p : 0]);
        double vchi = tested ? schi[p] : 0.0, vrss = tested ? srss[p] : 0.0;
        double vchi2 = vchi * vchi;
        // sum of r' over a point is not an invariant of the projection basis; sum_res carries sum ||r'||_1-free
        // moment 0 instead: we accumulate ||r'||^2 only (sum_res2) and leave sum_res = 0.
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          vchi += __shfl_xor_sync(0xffffffffu, vchi, o);
          vchi2 += __shfl_xor_sync(0xffffffffu, vchi2, o);
          vrss += __shfl_xor_sync(0xffffffffu, vrss, o);
        }
        s_chi += vchi, s_chi2 += vchi2, s_res2 += vrss;
        c_knn += __popc(__ballot_sync(0xffffffffu, stt == 1));
        c_plane += __popc(__ballot_sync(0xffffffffu, stt == 2));
        c_bad += __popc(__ballot_sync(0xffffffffu, stt == 3));
        c_chi += __popc(__ballot_sync(0xffffffffu, stt == 4));
        c_acc += __popc(__ballot_sync(0xffffffffu, stt == 0));
      }
    }
    __syncthreads();
  }
  Partial *out = &parts[blockIdx.x];
  if (e < ML_GSYM)
    out->G[e] = acc;
  else if (e < ML_GSYM + ML_G)
    out->b[e - ML_GSYM] = acc;
  else if (e == 192) {
    out->sum_chi = s_chi, out->sum_chi2 = s_chi2, out->sum_res = s_res, out->sum_res2 = s_res2;
    out->n_knn_fail = c_knn, out->n_plane_fail = c_plane, out->n_badmatch = c_bad, out->n_chi_rej = c_chi, out->n_acc = c_acc;
    out->pad = 0;
  }
}

// Sum the partials in a fixed order, expand to the state columns: HtH = T^T G T, Htr = T^T b. One block of 1024:
// thread (e, q) sums entry e over the partials p = q, q+4, ... with 8 loads in flight; the 4 groups are combined in order.
__global__ void __launch_bounds__(1024) k_finalize(const ScanState *__restrict__ st, const Partial *__restrict__ parts, int nparts,
                                                  const DevConsts *__restrict__ dc, LinsysOut *__restrict__ out) {
  __shared__ double G[ML_G][ML_G];
  __shared__ double bv[ML_G];
  __shared__ double GT[ML_G][ML_MAX_HX];
  __shared__ double grp[4][ML_GSYM + ML_G];
  const int hx = dc->c.total_hx;
  const int e = threadIdx.x & 255, q = threadIdx.x >> 8;
  if (e < ML_GSYM + ML_G) {
    double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const double *base = (e < ML_GSYM) ? &parts[0].G[e] : &parts[0].b[e - ML_GSYM];
    const size_t stride = sizeof(Partial) / sizeof(double);
    int p = q;
    for (; p + 28 < nparts; p += 32) {
#pragma unroll
      for (int u = 0; u < 8; u++)
        acc[u] += base[(size_t)(p + 4 * u) * stride];
    }
    for (int u = 0; p < nparts; p += 4, u++)
      acc[u] += base[(size_t)p * stride];
    grp[q][e] = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
  } else if (threadIdx.x == 200) {
    ml_counters cn;
    cn.n_points = st->n_scan;
    cn.n_knn_fail = cn.n_plane_fail = cn.n_badmatch = cn.n_chi_rej = cn.n_acc = 0;
    cn.sum_chi = cn.sum_chi2 = cn.sum_res = cn.sum_res2 = 0.0;
    for (int p = 0; p < nparts; p++) {
      cn.n_knn_fail += parts[p].n_knn_fail, cn.n_plane_fail += parts[p].n_plane_fail, cn.n_badmatch += parts[p].n_badmatch;
      cn.n_chi_rej += parts[p].n_chi_rej, cn.n_acc += parts[p].n_acc;
      cn.sum_chi += parts[p].sum_chi, cn.sum_chi2 += parts[p].sum_chi2, cn.sum_res += parts[p].sum_res, cn.sum_res2 += parts[p].sum_res2;
    }
    out->cnt = cn;
    out->st = *st;
  }
  __syncthreads();
  if (q == 0 && e < ML_GSYM + ML_G) {
    const double s = (grp[0][e] + grp[1][e]) + (grp[2][e] + grp[3][e]);
    if (e < ML_GSYM) {
      int r = 0, rem = e;
      while (rem >= ML_G - r) {
        rem -= ML_G - r;
        r++;
      }
      G[r][r + rem] = s;
      G[r + rem][r] = s;
    } else
      bv[e - ML_GSYM] = s;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < ML_G * hx; idx += blockDim.x) {
    const int r = idx / hx, j = idx - r * hx;
    double s = 0.0;
#pragma unroll
    for (int l = 0; l < ML_G; l++)
      s += G[r][l] * dc->T[l * ML_MAX_HX + j];
    GT[r][j] = s;
  }
  __syncthreads();
  // upper triangle computed once and mirrored: HtH is exactly symmetric
  for (int idx = threadIdx.x; idx < hx * hx; idx += blockDim.x) {
    const int a = idx / hx, b2 = idx - a * hx;
    if (a > b2)
      continue;
    double s = 0.0;
#pragma unroll
    for (int l = 0; l < ML_G; l++)
      s += dc->T[l * ML_MAX_HX + a] * GT[l][b2];
    out->HtH[idx] = s;
    out->HtH[b2 * hx + a] = s;
  }
  for (int a = threadIdx.x; a < hx; a += blockDim.x) {
    double s = 0.0;
#pragma unroll
    for (int l = 0; l < ML_G; l++)
      s += dc->T[l * ML_MAX_HX + a] * bv[l];
    out->Htr[a] = s;
  }
}

// parity export: h = T^T g for every point (n x hx)
__global__ void __launch_bounds__(256) k_expand_h(const ScanState *__restrict__ st, const PointRow *__restrict__ rows,
                                                 const DevConsts *__restrict__ dc, double *__restrict__ h_rows) {
  const int hx = dc->c.total_hx;
  const long long tot = (long long)st->n_scan * hx;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < tot; idx += (long long)gridDim.x * blockDim.x) {
    const int p = (int)(idx / hx), j = (int)(idx - (long long)p * hx);
    double s = 0.0;
    if (rows[p].status == 0 || rows[p].status == 4) {
#pragma unroll
      for (int l = 0; l < ML_G; l++)
        s += rows[p].g[l] * dc->T[l * ML_MAX_HX + j];
    }
    h_rows[idx] = s;
  }
}

} // namespace ml
