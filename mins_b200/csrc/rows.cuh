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

// quantile(chi_squared(dof), 0.95), the entries of UpdaterStatistics' chi_squared_table (UpdaterStatistics.cpp:31-33) the path can
// reach: the projected residual has map_ngbr_num + 1 - 3 rows (UpdaterLidar.cpp:301,362), map_ngbr_num in [3, ML_MAX_K]
__host__ __device__ constexpr double chi2_q95(int dof) {
  return dof == 1 ? 3.841458820694124 : dof == 2 ? 5.991464547107979 : dof == 3 ? 7.814727903251179 : dof == 4 ? 9.487729036781154
       : dof == 5 ? 11.070497693516351 : dof == 6 ? 12.591587243743977 : dof == 7 ? 14.067140449340169 : 15.50731305586545;
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

// T (18 x hx) and Q = T P T^T from the scan constants. One block of 1024; T and T P staged in shared memory, P read
// coalesced (thread index runs along P's rows).
// `c_in` and `P` may live in pinned HOST memory (the library's staging block): the kernel pulls the ~24 KB over PCIe itself, so
// the side stream carries no copy nodes at all.
__global__ void __launch_bounds__(1024) k_prep_consts(DevConsts *dc, const ml_scan_consts *__restrict__ c_in, const double *__restrict__ P) {
  __shared__ double sT[ML_G * ML_MAX_HX];
  __shared__ double sTP[ML_G * ML_MAX_HX];
  __shared__ double sP[ML_MAX_HX * ML_MAX_HX]; // 32 KB: P staged with every load in flight at once (one DRAM round trip)
  const int hx = c_in->total_hx;
  for (int e = threadIdx.x; e < hx * hx; e += blockDim.x)
    sP[e] = __ldg(&P[e]);
  {
    const int nw = (int)(sizeof(ml_scan_consts) / sizeof(int));
    const int *src = reinterpret_cast<const int *>(c_in);
    int *dst = reinterpret_cast<int *>(&dc->c);
    for (int e = threadIdx.x; e < nw; e += blockDim.x)
      dst[e] = __ldg(&src[e]);
  }
  __syncthreads(); // dc->c is complete and visible to the whole CTA from here on
  const ml_scan_consts &c = dc->c;
  for (int e = threadIdx.x; e < ML_G * ML_MAX_HX; e += blockDim.x)
    sT[e] = 0.0;
  __syncthreads();
  // columns of different blocks may coincide (I and A clones overlap): one thread per row adds its blocks in order
  if (threadIdx.x < ML_G) {
    const int r = threadIdx.x;
    double *Tr = sT + r * ML_MAX_HX;
    if (r < 6) { // I rotation rows 0-2, I position rows 3-5
      for (int b = 0; b < c.nI; b++)
        for (int cc = 0; cc < c.dTdxI[b].width; cc++)
          Tr[c.dTdxI[b].col + cc] += c.dTdxI[b].D[6 * r + cc];
    }
    if (r >= 3 && r < 6) { // ... minus A position rows (g_A,pos = -g_I,pos)
      for (int b = 0; b < c.nA; b++)
        for (int cc = 0; cc < c.dTdxA[b].width; cc++)
          Tr[c.dTdxA[b].col + cc] -= c.dTdxA[b].D[6 * r + cc];
    } else if (r >= 6 && r < 9) { // A rotation rows
      for (int b = 0; b < c.nA; b++)
        for (int cc = 0; cc < c.dTdxA[b].width; cc++)
          Tr[c.dTdxA[b].col + cc] += c.dTdxA[b].D[6 * (r - 6) + cc];
    } else if (r >= 9 && c.calib_col >= 0) {
      Tr[c.calib_col + (r - 9)] += 1.0;
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < ML_G * ML_MAX_HX; e += blockDim.x)
    dc->T[e] = sT[e];
  for (int e = threadIdx.x; e < ML_G * hx; e += blockDim.x) {
    const int r = e / hx, j = e - r * hx;
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    int l = 0;
    for (; l + 4 <= hx; l += 4) {
      s0 += sT[r * ML_MAX_HX + l] * sP[l * hx + j];
      s1 += sT[r * ML_MAX_HX + l + 1] * sP[(l + 1) * hx + j];
      s2 += sT[r * ML_MAX_HX + l + 2] * sP[(l + 2) * hx + j];
      s3 += sT[r * ML_MAX_HX + l + 3] * sP[(l + 3) * hx + j];
    }
    for (; l < hx; l++)
      s0 += sT[r * ML_MAX_HX + l] * sP[l * hx + j];
    sTP[r * ML_MAX_HX + j] = (s0 + s1) + (s2 + s3);
  }
  __syncthreads();
  for (int e = threadIdx.x; e < ML_G * ML_G; e += blockDim.x) {
    const int r = e / ML_G, q = e - r * ML_G;
    double s0 = 0, s1 = 0;
    int l = 0;
    for (; l + 2 <= hx; l += 2) {
      s0 += sTP[r * ML_MAX_HX + l] * sT[q * ML_MAX_HX + l];
      s1 += sTP[r * ML_MAX_HX + l + 1] * sT[q * ML_MAX_HX + l + 1];
    }
    for (; l < hx; l++)
      s0 += sTP[r * ML_MAX_HX + l] * sT[q * ML_MAX_HX + l];
    dc->Q[e] = s0 + s1;
  }
  __syncthreads();
  __threadfence_block();
  // packed upper triangle, off-diagonals doubled, for the per-point quadratic form g^T Q g
  for (int e = threadIdx.x; e < ML_GSYM; e += blockDim.x) {
    int r = 0, rem = e;
    while (rem >= ML_G - r) {
      rem -= ML_G - r;
      r++;
    }
    const int q = r + rem;
    dc->Qu[e] = (q == r) ? dc->Q[r * ML_G + r] : dc->Q[r * ML_G + q] + dc->Q[q * ML_G + r];
  }
}

// One scan point: neighbours -> plane -> gates -> g, w, vr, chi2 (see the header of this file). `c` and `sQ` live in
// shared memory; `out` arrives zeroed with status 1. out.status: 0 accepted, 1 kNN fail, 2 plane fail, 3 bad match,
// 4 chi2 reject.
template <int K, int NB>
__device__ __forceinline__ void point_row(const int i, const ml_scan_consts &c, const double *__restrict__ sQ,
                                          const float4 *__restrict__ scan_L, const float4 *__restrict__ map_pts,
                                          const int32_t *__restrict__ knn_idx, const uint8_t *__restrict__ knn_ok, PointRow &out) {
  if (!knn_ok[i]) {
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
    // (the A position block is -nG: folded into rows 3-5 of T, see ML_G)
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
      g[9 + a] = x1[a] - x2[a] + x3[a];
      g[12 + a] = n[a] - a3[a];
    }
    if (c.calib_col < 0) {
#pragma unroll
      for (int a = 9; a < ML_G; a++)
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
  // sQ = packed upper triangle of Q (ML_G x ML_G) by rows, off-diagonal entries doubled (Q is symmetric). Only the leading
  // 3*NB entries of g are non-zero (NB = 3 without extrinsic calibration: 45 terms; NB = 5: 120 terms).
  double s = 0.0;
  {
#pragma unroll
    for (int a = 0; a < 3 * NB; a++) {
      double acc = 0.0;
#pragma unroll
      for (int bq = a; bq < 3 * NB; bq++)
        acc += sQ[a * ML_G - (a * (a - 1)) / 2 + (bq - a)] * g[bq];
      s += g[a] * acc;
    }
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
  out.status = (chi2 < c.chi2_mult * chi2_q95(K - 2)) ? 0 : 4; // NaN -> reject (UpdaterStatistics.cpp:117)
}

// ------------------------------------------------------------------------------------------------
// Fused rows + reduction + finalize (one launch per scan). Warp-autonomous: no CTA-wide barrier in the main loop.
//  * the scan is cut into tiles of 32 points; tile j belongs to global warp j mod NW, with global warp id
//    gw = warp_in_cta * gridDim + cta (so a partial last round spreads over all SMs). A warp walks its tiles in order.
//  * phase 1 (32 lanes): lane computes point_row of its point and leaves g (ML_G = 15 folded numbers, 9 non-zero without
//    extrinsic calibration), w, v^T r in the warp's shared-memory slab (never in HBM unless the parity export `rows_out` is
//    requested); chi2 statistics by a fixed shuffle tree.
//  * phase 2: G = sum w g g^T is cut into 3x3 blocks; with NB = 3 (no calibration) or 5 block rows there are
//    NB(NB+1)/2 = 6 or 15 upper-triangle blocks ("roles"). The 32 lanes are split into 5 (or 2) groups of roles; group j
//    adds the points p = j, j + groups, ... of the tile, so one tile costs 7 (or 16) iterations of 7 shared loads and 9 to 12
//    FMAs instead of 32; diagonal blocks also carry b = sum g (v^T r). Running sums live in the warp's slab, one slot per lane.
//  * end: one barrier, CTA partial = sum over its warps and groups in fixed order -> parts[e][cta] (entry-major so the final
//    pass reads coalesced); the last CTA to finish (ticket) adds the partials in CTA order, expands HtH = T^T G T,
//    Htr = T^T b and writes the result block. Every sum has a fixed order for a given scan size and SM count: bit-reproducible.
// ------------------------------------------------------------------------------------------------
#ifndef ML_RR_BLOCK
#define ML_RR_BLOCK 512
#endif
constexpr int RR_BLOCK = ML_RR_BLOCK;
constexpr int RR_WARPS = RR_BLOCK / 32;
constexpr int RR_NE = ML_GSYM + ML_G; // 135
constexpr int RR_ROW = ML_G + 2;      // odd row stride of g in shared memory (bank-conflict-free stores)
constexpr int RR_RACC = 12;           // 9 block entries + 3 entries of b (diagonal blocks)
#define ML_NRED (RR_NE + 3)           // + sum_chi, sum_chi2, sum_res2
struct RRParts {
  double v[ML_NRED][ML_NPART_MAX];
  int c[5][ML_NPART_MAX];
};
// per-warp slab: g rows | w | vr | running sums (one slot per lane) | 4 statistics
constexpr int RR_SLAB = 32 * RR_ROW + 32 + 32 + 32 * RR_RACC + 4;
constexpr size_t RR_SMEM = sizeof(double) * ((size_t)RR_WARPS * RR_SLAB + ML_GSYM + 1) + sizeof(ml_scan_consts) + 64;

template <int NB> __device__ __forceinline__ int rr_role_of(int bi, int bj) { return bi * NB - (bi * (bi - 1)) / 2 + (bj - bi); }

template <int K, int NB>
__global__ void __launch_bounds__(RR_BLOCK, 1)
    k_rows_reduce(const ScanState *__restrict__ st, const float4 *__restrict__ scan_L, const float4 *__restrict__ map_pts,
                  const int32_t *__restrict__ knn_idx, const uint8_t *__restrict__ knn_ok, const DevConsts *__restrict__ dc,
                  PointRow *__restrict__ rows_out, RRParts *__restrict__ parts, unsigned int *__restrict__ ticket,
                  LinsysOut *__restrict__ out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *slabs = reinterpret_cast<double *>(smem_raw);
  double *sQ = slabs + (size_t)RR_WARPS * RR_SLAB; // packed Qu (171)
  ml_scan_consts *scp = reinterpret_cast<ml_scan_consts *>(sQ + ML_GSYM + 1);
  __shared__ int scnt[5];
  __shared__ unsigned int s_last;
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  pdl_enter();
  {
    const int nw = (int)(sizeof(ml_scan_consts) / sizeof(int));
    const int *src = reinterpret_cast<const int *>(&dc->c);
    int *dst = reinterpret_cast<int *>(scp);
    for (int e = t; e < nw; e += RR_BLOCK)
      dst[e] = src[e];
    for (int e = t; e < ML_GSYM; e += RR_BLOCK)
      sQ[e] = dc->Qu[e];
    if (t < 5)
      scnt[t] = 0;
  }
  double *sg = slabs + (size_t)wid * RR_SLAB; // [32][RR_ROW]
  double *sw = sg + 32 * RR_ROW;               // [32]
  double *svr = sw + 32;                       // [32]
  double *sacc = svr + 32;                     // [32 lanes][RR_RACC]
  double *sstat = sacc + 32 * RR_RACC;         // chi, chi^2, rss
  for (int e = lane; e < 32 * RR_RACC + 4; e += 32)
    sacc[e] = 0.0;
  __syncthreads();
  const ml_scan_consts &c = *scp;
  const int n = st->n_scan;
  const int nb = gridDim.x;
  const int NW = nb * RR_WARPS;
  const int gw = wid * nb + (int)blockIdx.x;
  const int ntiles = (n + 31) >> 5;
  // phase-2 role of this lane: group `grp` of points, block (bi, bj), bi <= bj
  constexpr int ROLES = NB * (NB + 1) / 2;
  constexpr int GROUPS = 32 / ROLES;
  const int grp = lane / ROLES, role = lane - grp * ROLES;
  const bool p2_lane = lane < ROLES * GROUPS;
  int bi = 0, bj = 0;
  {
    int r = 0, rem = role;
    while (rem >= NB - r) {
      rem -= NB - r;
      r++;
    }
    bi = r;
    bj = r + rem;
  }
  int c_st[5] = {0, 0, 0, 0, 0};

  for (int tile = gw; tile < ntiles; tile += NW) {
    const int i = (tile << 5) + lane;
    // ---- phase 1 ----
    PointRow pr;
#pragma unroll
    for (int a = 0; a < ML_G; a++)
      pr.g[a] = 0.0;
    pr.w = pr.vr = pr.chi2 = pr.res = pr.raw_n = pr.rss = 0.0;
    pr.plane[0] = pr.plane[1] = pr.plane[2] = pr.plane[3] = 0.0;
    pr.pad = 0;
    pr.status = 1;
    if (i < n) {
      point_row<K, NB>(i, c, sQ, scan_L, map_pts, knn_idx, knn_ok, pr);
      if (rows_out)
        rows_out[i] = pr;
    }
    const bool okp = i < n && pr.status == 0;
#pragma unroll
    for (int a = 0; a < ML_G; a++)
      sg[lane * RR_ROW + a] = okp ? pr.g[a] : 0.0;
    sw[lane] = okp ? pr.w : 0.0;
    svr[lane] = okp ? pr.vr : 0.0;
    {
      const bool tested = i < n && (pr.status == 0 || pr.status == 4) && (pr.chi2 == pr.chi2);
      double vchi = tested ? pr.chi2 : 0.0, vrss = tested ? pr.rss : 0.0;
      double vchi2 = vchi * vchi;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        vchi += __shfl_xor_sync(0xffffffffu, vchi, o);
        vchi2 += __shfl_xor_sync(0xffffffffu, vchi2, o);
        vrss += __shfl_xor_sync(0xffffffffu, vrss, o);
      }
      if (lane == 0) {
        sstat[0] += vchi;
        sstat[1] += vchi2;
        sstat[2] += vrss;
      }
#pragma unroll
      for (int q = 0; q < 5; q++)
        c_st[q] += __popc(__ballot_sync(0xffffffffu, i < n && pr.status == q));
    }
    const unsigned any_ok = __ballot_sync(0xffffffffu, okp);
    __syncwarp();
    // ---- phase 2 ----
    if (any_ok && p2_lane) {
      double a9[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      double b3[3] = {0, 0, 0};
      const double *gI = sg + 3 * bi, *gJ = sg + 3 * bj;
#pragma unroll 4
      for (int p = grp; p < 32; p += GROUPS) {
        const double w = sw[p];
        const double x0 = gI[p * RR_ROW] * w, x1 = gI[p * RR_ROW + 1] * w, x2 = gI[p * RR_ROW + 2] * w;
        const double y0 = gJ[p * RR_ROW], y1 = gJ[p * RR_ROW + 1], y2 = gJ[p * RR_ROW + 2];
        a9[0] += x0 * y0, a9[1] += x0 * y1, a9[2] += x0 * y2;
        a9[3] += x1 * y0, a9[4] += x1 * y1, a9[5] += x1 * y2;
        a9[6] += x2 * y0, a9[7] += x2 * y1, a9[8] += x2 * y2;
        if (bi == bj) {
          const double vr = svr[p];
          b3[0] += y0 * vr, b3[1] += y1 * vr, b3[2] += y2 * vr;
        }
      }
      double *my = sacc + lane * RR_RACC;
#pragma unroll
      for (int q = 0; q < 9; q++)
        my[q] += a9[q];
      if (bi == bj) {
        my[9] += b3[0], my[10] += b3[1], my[11] += b3[2];
      }
    }
    __syncwarp();
  }
  if (lane == 0) {
#pragma unroll
    for (int q = 0; q < 5; q++)
      if (c_st[q])
        atomicAdd(&scnt[q], c_st[q]);
  }
  __syncthreads();
  // ---- CTA partial: sum over the warps in warp order ----
  if (t < RR_NE) {
    int erole = -1, q = 0;
    if (t < ML_GSYM) {
      int r = 0, rem = t;
      while (rem >= ML_G - r) {
        rem -= ML_G - r;
        r++;
      }
      const int gi = r, gj = r + rem;
      if (gj < 3 * NB) {
        erole = rr_role_of<NB>(gi / 3, gj / 3);
        q = (gi % 3) * 3 + (gj % 3);
      }
    } else {
      const int a = t - ML_GSYM;
      if (a < 3 * NB) {
        erole = rr_role_of<NB>(a / 3, a / 3);
        q = 9 + a % 3;
      }
    }
    double v = 0.0;
    if (erole >= 0) {
      for (int w = 0; w < RR_WARPS; w++)
#pragma unroll
        for (int gq = 0; gq < GROUPS; gq++)
          v += slabs[(size_t)w * RR_SLAB + 32 * RR_ROW + 64 + (gq * ROLES + erole) * RR_RACC + q];
    }
    parts->v[t][blockIdx.x] = v;
  } else if (t < RR_NE + 3) {
    const int q = t - RR_NE;
    double v = 0.0;
    for (int w = 0; w < RR_WARPS; w++)
      v += slabs[(size_t)w * RR_SLAB + 32 * RR_ROW + 64 + 32 * RR_RACC + q];
    parts->v[t][blockIdx.x] = v;
  } else if (t >= 192 && t < 197) {
    parts->c[t - 192][blockIdx.x] = scnt[t - 192];
  }
  // ---- last CTA finalizes ----
  __threadfence();
  __syncthreads();
  if (t == 0) {
    const unsigned int prev = atomicAdd(ticket, 1u);
    s_last = (prev == (unsigned int)nb - 1u) ? 1u : 0u;
  }
  __syncthreads();
  if (!s_last)
    return;
  __threadfence();
  // shared-memory carve-up for the tail (the slabs are dead now)
  double *fsum = slabs;               // [ML_NRED]
  double *G = fsum + 256;             // [18][18]
  double *bv = G + ML_G * ML_G;       // [18]
  double *sT = bv + 32;               // [18][ML_MAX_HX]
  double *GT = sT + ML_G * ML_MAX_HX; // [18][ML_MAX_HX]
  const int hx = c.total_hx;
  const volatile RRParts *vp = parts; // written by other CTAs: do not serve from L1
  for (int q = wid; q < ML_NRED; q += RR_WARPS) {
    double v = 0.0;
    for (int bq = lane; bq < nb; bq += 32)
      v += vp->v[q][bq];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0)
      fsum[q] = v;
  }
  if (wid < 5) { // integer counters
    int v = 0;
    for (int bq = lane; bq < nb; bq += 32)
      v += vp->c[wid][bq];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0)
      scnt[wid] = v;
  }
  for (int q = t; q < ML_G * ML_MAX_HX; q += RR_BLOCK)
    sT[q] = dc->T[q];
  __syncthreads();
  if (t < ML_GSYM) {
    int r = 0, rem = t;
    while (rem >= ML_G - r) {
      rem -= ML_G - r;
      r++;
    }
    G[r * ML_G + r + rem] = fsum[t];
    G[(r + rem) * ML_G + r] = fsum[t];
  } else if (t < RR_NE) {
    bv[t - ML_GSYM] = fsum[t];
  } else if (t == RR_NE) {
    ml_counters cn;
    cn.n_points = n;
    cn.n_acc = scnt[0], cn.n_knn_fail = scnt[1], cn.n_plane_fail = scnt[2], cn.n_badmatch = scnt[3], cn.n_chi_rej = scnt[4];
    cn.sum_chi = fsum[RR_NE + 0], cn.sum_chi2 = fsum[RR_NE + 1], cn.sum_res = 0.0, cn.sum_res2 = fsum[RR_NE + 2];
    out->cnt = cn;
    out->st = *st;
    *ticket = 0u; // ready for the next scan
  }
  __syncthreads();
  for (int idx = t; idx < ML_G * hx; idx += RR_BLOCK) {
    const int r = idx / hx, j = idx - r * hx;
    double s = 0.0;
#pragma unroll
    for (int l = 0; l < ML_G; l++)
      s += G[r * ML_G + l] * sT[l * ML_MAX_HX + j];
    GT[r * ML_MAX_HX + j] = s;
  }
  __syncthreads();
  // upper triangle computed once and mirrored: HtH is exactly symmetric. Staged in shared memory and written out in
  // full 128-byte lines: `out` is pinned HOST memory (the result crosses PCIe as ~150 posted writes, no D2H copy node).
  double *Hs = GT + ML_G * ML_MAX_HX; // [hx][hx] + [hx]
  for (int idx = t; idx < hx * hx; idx += RR_BLOCK) {
    const int a = idx / hx, b2 = idx - a * hx;
    if (a > b2)
      continue;
    double s = 0.0;
#pragma unroll
    for (int l = 0; l < ML_G; l++)
      s += sT[l * ML_MAX_HX + a] * GT[l * ML_MAX_HX + b2];
    Hs[idx] = s;
    Hs[b2 * hx + a] = s;
  }
  for (int a = t; a < hx; a += RR_BLOCK) {
    double s = 0.0;
#pragma unroll
    for (int l = 0; l < ML_G; l++)
      s += sT[l * ML_MAX_HX + a] * bv[l];
    Hs[hx * hx + a] = s;
  }
  __syncthreads();
  for (int idx = t; idx < hx * hx; idx += RR_BLOCK)
    out->HtH[idx] = Hs[idx];
  for (int a = t; a < hx; a += RR_BLOCK)
    out->Htr[a] = Hs[hx * hx + a];
  __threadfence_system();
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
