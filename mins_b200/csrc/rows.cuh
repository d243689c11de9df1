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

// sqrt(x) for x >= 0 as x * rsqrt(x): MUFU.RSQ64H + two Newton steps, no slow-path call (<= 2 ulp; 0 -> 0)
__device__ __forceinline__ double fast_sqrt(double x) { return x > 0.0 ? x * rsqrt(x) : 0.0; }

// eigenvalues (min, max) of a symmetric 3x3 matrix [a00 a01 a02; . a11 a12; . . a22]: the trigonometric closed form
//   lambda = q + 2 p cos(phi + 2 pi k / 3),  phi = acos(r) / 3,  r = det((A - q I) / p) / 2,
// evaluated WITHOUT acos / cos (double-precision acos + 2 cos cost ~570 instructions, a third of the whole point):
// c = cos(phi) in [1/2, 1] is the largest root of 4 c^3 - 3 c = r. Start above it at c0 = 1/2 + sqrt((1 + r) / 6) (exact to
// first order at the double root r = -1, 1.08 at r = 1), three Newton steps in fp32 and two in fp64 (monotone from above: the
// cubic is convex and increasing there; error < 2e-12 at the ill-conditioned double root, where acos is no better, ~1e-16
// elsewhere). Then cos(phi + 2 pi / 3) = -(c + sqrt(3) sqrt(1 - c^2)) / 2.
__device__ __forceinline__ void sym3_eig_minmax(double a00, double a01, double a02, double a11, double a12, double a22, double *lmin,
                                                double *lmax) {
  const double p1 = fma(a01, a01, fma(a02, a02, a12 * a12));
  const double q = (a00 + a11 + a22) * (1.0 / 3.0);
  const double b00 = a00 - q, b11 = a11 - q, b22 = a22 - q;
  const double p2 = fma(b00, b00, fma(b11, b11, fma(b22, b22, 2.0 * p1)));
  if (!(p2 > 0.0)) {
    *lmin = *lmax = q;
    return;
  }
  const double ip = rsqrt(p2 * (1.0 / 6.0));
  const double p = p2 * (1.0 / 6.0) * ip;
  const double det = fma(b00, fma(b11, b22, -a12 * a12), fma(-a01, fma(a01, b22, -a12 * a02), a02 * fma(a01, a12, -b11 * a02)));
  double r = 0.5 * det * ip * ip * ip;
  r = fmin(1.0, fmax(-1.0, r));
  float cf = 0.5f + sqrtf((1.0f + (float)r) * (1.0f / 6.0f));
  const float rf = (float)r;
#pragma unroll
  for (int it = 0; it < 3; it++) {
    const float c2 = cf * cf;
    cf = cf - __fdividef(fmaf(fmaf(4.0f, c2, -3.0f), cf, -rf), fmaf(12.0f, c2, -3.0f) + 1e-30f);
  }
  double c = fmax(0.5, (double)cf);
#pragma unroll
  for (int it = 0; it < 2; it++) {
    const double c2 = c * c;
    const double fp = fma(12.0, c2, -3.0);
    const double f = fma(fma(4.0, c2, -3.0), c, -r);
    c = fp > 1e-300 ? c - f / fp : c;
  }
  c = fmin(1.0, fmax(0.5, c));
  const double sn = fast_sqrt(fma(-c, c, 1.0));
  *lmax = fma(2.0 * p, c, q);
  *lmin = q - p * fma(1.7320508075688772, sn, c);
}

// ------------------------------------------------------------------------------------------------
// Per-scan constants of the point loop in the form the kernel consumes them, derived on the host from ml_scan_consts
// (make_row_consts, api.cu) and passed BY VALUE as a __grid_constant__ kernel parameter: parameters live in the constant
// bank, so every use is an instruction operand - no load, no register, no shared-memory staging.
//   w  = RLtoI (pfinL - pIinL) = RLtoI pfinL + pLinI                (UpdaterLidar.cpp:320)
//   nI = (RGtoM RItoG)^T n,  nG = RGtoM^T n                         (:320-321)
//   u  = RGtoA (pIinG + RItoG pLinI - pAinG + RLtoG pfinL) = U0 + U1 pfinL   (:322)
//   c1 = RGtoM (pIinG - pAinG)                                      (:325)
// ------------------------------------------------------------------------------------------------
struct RowConsts {
  double RLtoM_est[9], pLinM_est[3], RLtoM_fej[9], pLinM_fej[3];
  double RLtoI[9], pLinI[3];
  double RItoM[9], RGtoM[9], RItoL[9];
  double U1[9], U0[3];
  double c1[3], pIinL[3];
  double raw_noise, map_noise;
  double condi2;       // plane_max_condi^2
  double p2pd;         // plane_max_p2pd
  double badmatch;     // 3 plane_max_p2pd (UpdaterLidar.cpp:296)
  double ie_o, ie_p;
  double chi2_thr;     // chi2_mult * quantile(chi_squared(K - 2), 0.95)
  int intr;            // !at_clone_m || !at_clone_a (UpdaterLidar.cpp:332)
  int total_hx;
  // Q = T P T^T restricted to the 3 NB directions that can be non-zero, upper triangle by rows with row stride 3 NB and the
  // off-diagonal entries doubled: h^T P h = g^T Q g = sum_a g_a sum_{b >= a} Qu[a][b] g_b (host: make_row_consts)
  double Qu[ML_G * ML_G];
};

// One scan point: k neighbour coordinates -> plane -> gates -> g (UN-whitened, 3 NB numbers), weights, chi2.
//   Gram route: G = A^T A (3x3, uncentred, as the reference's A), condition gate from its extreme eigenvalues
//   (cond(A)^2 = lmax / lmin, LidarHelper.cpp:305-312), plane abc = -d G^-1 A^T 1 by Cholesky G = L L^T (LidarHelper.cpp:315:
//   the least-squares solution colPivHouseholderQr returns; the gate bounds cond(G) by plane_max_condi^2 = 4e4, so the normal
//   equations lose < 5 of 16 digits: 1e-11 against the 1e-9 contract). The neighbour residuals e_j = abc p_j + d serve the
//   sanity gate (:319-325), the residual rows (UpdaterLidar.cpp:305-310) and - being the least-squares residual of that very
//   fit, orthogonal to the columns of A - the neighbour part of ||r'||^2 directly: RSS_nb = sum e_j^2 / map_n^2, and the
//   deleted residual of the scan-point row is its plain whitened residual e5 = res5 / raw_n. Leverage of the scan-point row:
//   t = (map_n / raw_n)^2 || L^-1 pfinM_fej ||^2. Then ||v||^2 = 1/(1+t), v^T r' = e5/(1+t), ||r'||^2 = RSS_nb + e5^2/(1+t),
//   chi2 = ||r'||^2 - s (v^T r')^2 / (1 + s ||v||^2), s = h^T P h = g^T Q g / raw_n^2 (header of this file).
// Returns the status (0 accepted, 2 plane fail, 3 bad match, 4 chi2 reject); the caller has checked knn_ok.
// Outputs: g[3 NB] un-whitened; W = ||v||^2 / raw_n^2 (weight of g g^T), V = v^T r' / raw_n (weight of g in H^T r).
struct RowOut {
  double W, V, chi2, rss, raw_n, res5;
  double n[3], d;
  double sum_rp; // STATS: sum of the K - 2 projected residuals r' in the reference's own nullspace basis
};
// Eigen::JacobiRotation<double>::makeGivens(p, q), real case: G = [c s; -s c] with G^* [p; q] = [r; 0]
__device__ __forceinline__ void make_givens(double p, double q, double &c, double &s) {
  if (q == 0.0) {
    c = p < 0 ? -1.0 : 1.0, s = 0.0;
  } else if (p == 0.0) {
    c = 0.0, s = q < 0 ? 1.0 : -1.0;
  } else if (fabs(p) > fabs(q)) {
    const double t = q / p;
    double u = sqrt(1.0 + t * t);
    u = p < 0 ? -u : u;
    c = 1.0 / u, s = -t * c;
  } else {
    const double t = p / q;
    double u = sqrt(1.0 + t * t);
    u = q < 0 ? -u : u;
    s = -1.0 / u, c = -t * s;
  }
}
// STATS (residual statistics switched on, ml_set_stats): also the FIRST moment of the projected residuals, which - unlike
// ||r'||^2 and chi2 - depends on the basis of the left nullspace, so the reference's own Givens sequence
// (StateHelper::Givens_Rotation, StateHelper.cpp:660-680: columns left to right, rows bottom up) is run on the whitened
// (K + 1) x 3 H_f and the residual; costs about a third of the point on top.
template <int K, int NB, bool STATS>
__device__ __forceinline__ int point_row(const RowConsts &rc, const float4 pl4, const float4 (&nb)[K], double *__restrict__ g,
                                         RowOut &o) {
  // ---- neighbours: Gram matrix, column sums, mean range (LidarHelper.cpp:293-302) ----
  float px[K], py[K], pz[K];
  double g00 = 0, g01 = 0, g02 = 0, g11 = 0, g12 = 0, g22 = 0, sx = 0, sy = 0, sz = 0, dsum = 0;
#pragma unroll
  for (int j = 0; j < K; j++) {
    const float4 p = nb[j];
    px[j] = p.x, py[j] = p.y, pz[j] = p.z;
    const double x = (double)p.x, y = (double)p.y, z = (double)p.z;
    g00 = fma(x, x, g00), g01 = fma(x, y, g01), g02 = fma(x, z, g02);
    g11 = fma(y, y, g11), g12 = fma(y, z, g12), g22 = fma(z, z, g22);
    sx += x, sy += y, sz += z;
    dsum += fast_sqrt(fma(x, x, fma(y, y, z * z)));
  }
  const double d = dsum * (1.0 / (double)K);
  // ---- condition number gate (LidarHelper.cpp:305-312): cond(A)^2 = lmax/lmin of A^T A ----
  {
    double lmin, lmax;
    sym3_eig_minmax(g00, g01, g02, g11, g12, g22, &lmin, &lmax);
    // condA > max  <=>  lmax > max^2 lmin. lmin <= 0 after rounding = singular: sigma_max / sigma_min is huge or inf in the
    // reference (> max: rejected). Its 0/0 = NaN case needs every neighbour at the origin, which a voxel-thinned map cannot hold.
    if (!(lmin > 0.0) || lmax > rc.condi2 * lmin)
      return 2;
  }
  // ---- Cholesky G = L L^T (inverse pivots i0, i1, i2), abc = -d G^-1 s ----
  const double i0 = rsqrt(g00);
  const double l10 = g01 * i0, l20 = g02 * i0;
  const double i1 = rsqrt(fma(-l10, l10, g11));
  const double l21 = fma(-l20, l10, g12) * i1;
  const double i2 = rsqrt(fma(-l21, l21, fma(-l20, l20, g22)));
  double n0, n1, n2;
  {
    const double z0 = -d * sx * i0;
    const double z1 = fma(-l10, z0, -d * sy) * i1;
    const double z2 = fma(-l21, z1, fma(-l20, z0, -d * sz)) * i2;
    n2 = z2 * i2;
    n1 = fma(-l21, n2, z1) * i1;
    n0 = fma(-l20, n2, fma(-l10, n1, z0)) * i0;
  }
  const double nn = fast_sqrt(fma(n0, n0, fma(n1, n1, n2 * n2)));
  // ---- sanity check (LidarHelper.cpp:319-325) and the neighbours' residuals (UpdaterLidar.cpp:305-310) in one pass ----
  double rss_nb = 0.0;
  {
    const double lim = rc.p2pd * nn;
    bool bad = !(nn == nn);
#pragma unroll
    for (int j = 0; j < K; j++) {
      const double e = fma(n0, (double)px[j], fma(n1, (double)py[j], fma(n2, (double)pz[j], d)));
      bad = bad || (fabs(e) > lim);
      rss_nb = fma(e, e, rss_nb);
    }
    if (bad)
      return 2;
  }
  o.n[0] = n0, o.n[1] = n1, o.n[2] = n2, o.d = d;
  // ---- scan point in the map frame, est and fej (UpdaterLidar.cpp:281-283); bad match gate + residual (:296,:313) ----
  const double qx = (double)pl4.x, qy = (double)pl4.y, qz = (double)pl4.z;
  double res5;
  {
    const double mx = fma(rc.RLtoM_est[0], qx, fma(rc.RLtoM_est[1], qy, fma(rc.RLtoM_est[2], qz, rc.pLinM_est[0])));
    const double my = fma(rc.RLtoM_est[3], qx, fma(rc.RLtoM_est[4], qy, fma(rc.RLtoM_est[5], qz, rc.pLinM_est[1])));
    const double mz = fma(rc.RLtoM_est[6], qx, fma(rc.RLtoM_est[7], qy, fma(rc.RLtoM_est[8], qz, rc.pLinM_est[2])));
    res5 = -fma(n0, mx, fma(n1, my, n2 * mz)) - d;
  }
  if (res5 > rc.badmatch)
    return 3;
  // ---- leverage of the scan-point row: || L^-1 pfinM_fej ||^2 ----
  double zz, fej[3];
  {
    const double fx = fma(rc.RLtoM_fej[0], qx, fma(rc.RLtoM_fej[1], qy, fma(rc.RLtoM_fej[2], qz, rc.pLinM_fej[0])));
    const double fy = fma(rc.RLtoM_fej[3], qx, fma(rc.RLtoM_fej[4], qy, fma(rc.RLtoM_fej[5], qz, rc.pLinM_fej[1])));
    const double fz = fma(rc.RLtoM_fej[6], qx, fma(rc.RLtoM_fej[7], qy, fma(rc.RLtoM_fej[8], qz, rc.pLinM_fej[2])));
    const double y0 = fx * i0;
    const double y1 = fma(-l10, y0, fy) * i1;
    const double y2 = fma(-l21, y1, fma(-l20, y0, fz)) * i2;
    zz = fma(y0, y0, fma(y1, y1, y2 * y2));
    fej[0] = fx, fej[1] = fy, fej[2] = fz;
  }
  // ---- g = n [dpfinM_dI | dpfinM_dA | dpfinM_dcalib], folded (UpdaterLidar.cpp:320-329; ML_G in common.cuh) ----
  {
    // -n^T RGtoM RItoG skew(w) = (w x nI)^T, w = RLtoI (pfinL - pIinL)
    const double wx = fma(rc.RLtoI[0], qx, fma(rc.RLtoI[1], qy, fma(rc.RLtoI[2], qz, rc.pLinI[0])));
    const double wy = fma(rc.RLtoI[3], qx, fma(rc.RLtoI[4], qy, fma(rc.RLtoI[5], qz, rc.pLinI[1])));
    const double wz = fma(rc.RLtoI[6], qx, fma(rc.RLtoI[7], qy, fma(rc.RLtoI[8], qz, rc.pLinI[2])));
    const double ix = fma(rc.RItoM[0], n0, fma(rc.RItoM[3], n1, rc.RItoM[6] * n2));
    const double iy = fma(rc.RItoM[1], n0, fma(rc.RItoM[4], n1, rc.RItoM[7] * n2));
    const double iz = fma(rc.RItoM[2], n0, fma(rc.RItoM[5], n1, rc.RItoM[8] * n2));
    g[0] = fma(wy, iz, -wz * iy), g[1] = fma(wz, ix, -wx * iz), g[2] = fma(wx, iy, -wy * ix);
    g[3] = fma(rc.RGtoM[0], n0, fma(rc.RGtoM[3], n1, rc.RGtoM[6] * n2));
    g[4] = fma(rc.RGtoM[1], n0, fma(rc.RGtoM[4], n1, rc.RGtoM[7] * n2));
    g[5] = fma(rc.RGtoM[2], n0, fma(rc.RGtoM[5], n1, rc.RGtoM[8] * n2));
    // n^T RItoL skew(u) = (a2 x u)^T, a2 = RItoL^T n
    const double ax = fma(rc.RItoL[0], n0, fma(rc.RItoL[3], n1, rc.RItoL[6] * n2));
    const double ay = fma(rc.RItoL[1], n0, fma(rc.RItoL[4], n1, rc.RItoL[7] * n2));
    const double az = fma(rc.RItoL[2], n0, fma(rc.RItoL[5], n1, rc.RItoL[8] * n2));
    const double ux = fma(rc.U1[0], qx, fma(rc.U1[1], qy, fma(rc.U1[2], qz, rc.U0[0])));
    const double uy = fma(rc.U1[3], qx, fma(rc.U1[4], qy, fma(rc.U1[5], qz, rc.U0[1])));
    const double uz = fma(rc.U1[6], qx, fma(rc.U1[7], qy, fma(rc.U1[8], qz, rc.U0[2])));
    g[6] = fma(ay, uz, -az * uy), g[7] = fma(az, ux, -ax * uz), g[8] = fma(ax, uy, -ay * ux);
    if (NB == 5) {
      // n^T (skew(c1) - skew(c2) + RLtoM skew(c3)) = n x (c1 - c2) + a3 x c3, c3 = pIinL - pfinL, c2 = RLtoM c3, a3 = RLtoM^T n
      const double c3x = rc.pIinL[0] - qx, c3y = rc.pIinL[1] - qy, c3z = rc.pIinL[2] - qz;
      const double ex = rc.c1[0] - fma(rc.RLtoM_fej[0], c3x, fma(rc.RLtoM_fej[1], c3y, rc.RLtoM_fej[2] * c3z));
      const double ey = rc.c1[1] - fma(rc.RLtoM_fej[3], c3x, fma(rc.RLtoM_fej[4], c3y, rc.RLtoM_fej[5] * c3z));
      const double ez = rc.c1[2] - fma(rc.RLtoM_fej[6], c3x, fma(rc.RLtoM_fej[7], c3y, rc.RLtoM_fej[8] * c3z));
      const double tx = fma(rc.RLtoM_fej[0], n0, fma(rc.RLtoM_fej[3], n1, rc.RLtoM_fej[6] * n2));
      const double ty = fma(rc.RLtoM_fej[1], n0, fma(rc.RLtoM_fej[4], n1, rc.RLtoM_fej[7] * n2));
      const double tz = fma(rc.RLtoM_fej[2], n0, fma(rc.RLtoM_fej[5], n1, rc.RLtoM_fej[8] * n2));
      g[9] = fma(n1, ez, -n2 * ey) + fma(ty, c3z, -tz * c3y);
      g[10] = fma(n2, ex, -n0 * ez) + fma(tz, c3x, -tx * c3z);
      g[11] = fma(n0, ey, -n1 * ex) + fma(tx, c3y, -ty * c3x);
      g[12] = n0 - tx, g[13] = n1 - ty, g[14] = n2 - tz;
    }
  }
  // ---- interpolation-error inflation (UpdaterLidar.cpp:332-337) and whitening scales (:353-354) ----
  double intr_std = 0.0;
  if (rc.intr)
    intr_std = fast_sqrt(fma(rc.ie_o, fma(g[0], g[0], fma(g[1], g[1], g[2] * g[2])), rc.ie_p * fma(g[3], g[3], fma(g[4], g[4], g[5] * g[5]))));
  const double raw_n = fma(nn, rc.raw_noise, intr_std);
  const double map_n = fma(nn, rc.map_noise, intr_std);
  const double inv = 1.0 / (raw_n * map_n); // one division for both reciprocals
  const double ir = inv * map_n, im = inv * raw_n;
  const double ratio = map_n * ir;
  const double wv = 1.0 / fma(ratio * ratio, zz, 1.0); // ||v||^2 = 1 / (1 + t)
  const double e5 = res5 * ir;
  const double vr = e5 * wv;
  const double rss = fma(rss_nb * im, im, e5 * e5 * wv);
  // ---- s = h^T P h = g^T Q g / raw_n^2 (rc.Qu: kernel parameter = constant-bank operands of the FMAs) ----
  double s = 0.0;
#pragma unroll
  for (int a = 0; a < 3 * NB; a++) {
    double acc = 0.0;
#pragma unroll
    for (int b = a; b < 3 * NB; b++)
      acc = fma(rc.Qu[a * (3 * NB) + b], g[b], acc);
    s = fma(g[a], acc, s);
  }
  s *= ir * ir;
  const double chi2 = rss - s * vr * vr / fma(s, wv, 1.0);
  if (STATS) {
    double A[K + 1][3], r[K + 1];
#pragma unroll
    for (int j = 0; j < K; j++) {
      A[j][0] = (double)px[j] * im, A[j][1] = (double)py[j] * im, A[j][2] = (double)pz[j] * im;
      r[j] = -fma(n0, (double)px[j], fma(n1, (double)py[j], fma(n2, (double)pz[j], d))) * im;
    }
    A[K][0] = fej[0] * ir, A[K][1] = fej[1] * ir, A[K][2] = fej[2] * ir;
    r[K] = res5 * ir;
#pragma unroll
    for (int cn = 0; cn < 3; cn++) {
#pragma unroll
      for (int m = K; m > cn; m--) {
        if (A[m][cn] == 0.0)
          continue;
        double c, s;
        make_givens(A[m - 1][cn], A[m][cn], c, s);
#pragma unroll
        for (int q = 0; q < 3; q++) {
          if (q < cn)
            continue;
          const double x = A[m - 1][q], y = A[m][q];
          A[m - 1][q] = c * x - s * y, A[m][q] = s * x + c * y;
        }
        A[m][cn] = 0.0;
        const double x = r[m - 1], y = r[m];
        r[m - 1] = c * x - s * y, r[m] = s * x + c * y;
      }
    }
    double sr = 0.0;
#pragma unroll
    for (int m = 3; m <= K; m++)
      sr += r[m];
    o.sum_rp = sr;
  }
  o.W = wv * ir * ir;
  o.V = vr * ir;
  o.chi2 = chi2;
  o.rss = rss;
  o.raw_n = raw_n;
  o.res5 = res5;
  return (chi2 < rc.chi2_thr) ? 0 : 4; // NaN -> reject (UpdaterStatistics.cpp:117)
}

// ------------------------------------------------------------------------------------------------
// Fused rows + reduction + finalize (one launch per scan). Warp-autonomous: no CTA-wide barrier in the main loop.
//  * the scan is cut into tiles of 32 points; tile j belongs to global warp j mod NW, with global warp id
//    gw = warp_in_cta * gridDim + cta (so a partial last round spreads over all SMs). A warp walks its tiles in order.
//    The host picks the block size so that the number of tile rounds per warp is whole or nearly so (launch_rows_nb).
//  * phase 1 (32 lanes): lane computes point_row of its point from the k neighbour coordinates the kNN kernel left in `nbr`
//    (coalesced: 16 k bytes per point, scan order) and leaves g (3 NB numbers, un-whitened), W, V in the warp's shared-memory
//    slab (never in HBM unless the parity export `rows_out` is requested); chi2 statistics by a fixed shuffle tree.
//  * phase 2: G = sum W g g^T is cut into 3x3 blocks; with NB = 3 (no calibration) or 5 block rows there are
//    NB(NB+1)/2 = 6 or 15 upper-triangle blocks ("roles"). The 32 lanes are split into 5 (or 2) groups of roles; group j
//    adds the points p = j, j + groups, ... of the tile, so one tile costs 7 (or 16) iterations of 7 shared loads and 9 to 12
//    FMAs instead of 32; diagonal blocks also carry b = sum g V. Running sums live in the warp's slab, one slot per lane.
//  * end: one barrier, CTA partial = sum over its warps and groups in fixed order -> parts[e][cta] (entry-major so the final
//    pass reads coalesced); the last CTA to finish (ticket) adds the partials in CTA order, expands HtH = T^T G T,
//    Htr = T^T b and writes the result block. Every sum has a fixed order for a given scan size and launch shape: bit-reproducible.
// ------------------------------------------------------------------------------------------------
constexpr int RR_BLOCK = 512, RR_BLOCK_MID = 896, RR_BLOCK_BIG = 1024; // 16, 28 or 32 warps per SM (one CTA per SM either way)
constexpr int RR_RACC = 12;           // 9 block entries + 3 entries of b (diagonal blocks)
// reduced entries for NB block rows: upper triangle of G (NG (NG + 1) / 2), b (NG), sum_chi, sum_chi2, sum_res2, sum_res
#define ML_NRED (ML_GSYM + ML_G + 4)
// What one CTA returns: its reduced system in the folded pose space (G = sum W g g^T upper triangle by rows, b = sum g V),
// the chi2 statistics and the status counters. Lives in PINNED HOST memory: the CTA's last act is ~0.5 KB of posted PCIe
// writes; no ticket, no fence, no serial tail on the device. The library's host tail (rows_host_tail, api.cu) adds the
// blocks in CTA order - a fixed order, so the result is bit-reproducible - and applies the chain rule HtH = T^T G T.
struct RRPart {
  double v[ML_NRED];
  int c[5];
  int n_scan; // the device-side size of the scan (k_scan_tile: only the CTA that owned the last tile knows it, the others write -1)
};
// k_scan_tile's parts are ARMED by the host before the launch: every word the CTA will write holds a pattern no result can be (a
// quiet NaN with a payload; INT_MIN for the counters), so the host can consume a part word by word as the posted writes land.
constexpr unsigned long long RR_ARMED_F64 = 0x7ff8dead5eed0001ull;
constexpr int RR_ARMED_I32 = (int)0x80000000;
// per-warp slab: g rows (odd stride: conflict-free stores) | W | V | running sums (one slot per lane) | statistics
template <int NB> struct RRLayout {
  static constexpr int NG = 3 * NB;
  static constexpr int NSYM = NG * (NG + 1) / 2, NE = NSYM + NG, NRED = NE + 4;
  static constexpr int ROW = NG + 2;
  static constexpr int SLAB = 32 * ROW + 32 + 32 + 32 * RR_RACC + 8; // + chi, chi^2, rss | 5 status counters (as int)
  static constexpr size_t smem(int warps) { return sizeof(double) * (size_t)(warps * SLAB); }
};

template <int NB> __device__ __forceinline__ int rr_role_of(int bi, int bj) { return bi * NB - (bi * (bi - 1)) / 2 + (bj - bi); }

// ------------------------------------------------------------------------------------------------
// Fused rows + reduction (one launch per scan). Warp-autonomous: no CTA-wide barrier in the main loop.
//  * the scan is cut into tiles of 32 points; round r covers the tiles [r nb, (r + 1) nb) (nb = CTAs), one tile per CTA, rotated
//    by r; warp w of a CTA walks the rounds w, w + WARPS, ... The host picks the block size so that the rounds per warp are
//    whole or nearly so (launch_rows_nb).
//  * phase 1 (32 lanes): lane computes point_row of its point from the k neighbour coordinates the kNN kernel left in `nbr`
//    (coalesced: 16 k bytes per point, scan order) and leaves g (3 NB numbers, un-whitened), W, V in the warp's shared-memory
//    slab (never in HBM unless the parity export is requested); chi2 statistics by a fixed shuffle tree.
//  * phase 2: G = sum W g g^T is cut into 3x3 blocks; with NB = 3 (no calibration) or 5 block rows there are
//    NB(NB+1)/2 = 6 or 15 upper-triangle blocks ("roles"). The 32 lanes are split into 5 (or 2) groups of roles; group j
//    adds the points p = j, j + groups, ... of the tile, so one tile costs 7 (or 16) iterations of 7 shared loads and 9 to 12
//    FMAs instead of 32; diagonal blocks also carry b = sum g V. Running sums live in the warp's slab, one slot per lane.
//  * end: one barrier, CTA partial = sum over its warps and groups in fixed order -> this CTA's RRPart in pinned host memory.
//    Every sum has a fixed order for a given scan size and launch shape: bit-reproducible.
// EXPORT: also write the per-point rows (parity export, ml_point_rows); the timed instantiation carries none of that state.
// ------------------------------------------------------------------------------------------------
template <int K, int NB, int BLOCK, bool EXPORT>
__global__ void __launch_bounds__(BLOCK, 1)
    k_rows_reduce(const __grid_constant__ RowConsts rc, const ScanState *__restrict__ st, int n_ub, const float4 *__restrict__ scan_L,
                  const float4 *__restrict__ nbr, PointRow *__restrict__ rows_out, RRPart *__restrict__ parts /* pinned host */,
                  unsigned long long *__restrict__ tl /* debug timeline (ml_debug_timeline) or NULL */) {
  using LY = RRLayout<NB>;
  tl_mark(tl, 0);
  constexpr int NG = LY::NG, ROW = LY::ROW, SLAB = LY::SLAB, WARPS = BLOCK / 32, NSYM = LY::NSYM, NE = LY::NE, NRED = LY::NRED;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *slabs = reinterpret_cast<double *>(smem_raw);
  __shared__ int scnt[5];
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  const int nb = gridDim.x;
  // Tile -> warp: round r = wid, wid + WARPS, ... covers the nb tiles [r nb, (r + 1) nb); inside a round the CTAs are rotated
  // by r, so a CTA's tiles do not share a residue mod nb. A spinning LiDAR's scan is periodic in the point index (rings x
  // azimuths: 32 consecutive points are the lower or the upper half of a 64-beam column) and a plain stride nb gave the even
  // CTAs all the ground returns and the odd ones the sky: 11 us against 4.6 us of tile loop on neighbouring SMs.
  const auto tile_of = [&](int r) { return r * nb + (((int)blockIdx.x + nb - r % nb) % nb); };
  pdl_launch();
  if (t < 5)
    scnt[t] = 0;
  double *sg = slabs + (size_t)wid * SLAB; // [32][ROW]
  double *sw = sg + 32 * ROW;              // [32]
  double *sv = sw + 32;                    // [32]
  double *sacc = sv + 32;                  // [32 lanes][RR_RACC]
  double *sstat = sacc + 32 * RR_RACC;     // chi, chi^2, rss
  int *scst = reinterpret_cast<int *>(sstat + 3); // 5 status counters of this warp
  for (int e = lane; e < 32 * RR_RACC + 8; e += 32)
    sacc[e] = 0.0;
  pdl_wait();
  tl_mark(tl, 1);
  // the inputs of this warp's first tile and the scan size are requested together (n_ub = host-side bound of the scan size;
  // what lies past the device-side count is stale but readable, and discarded below)
  float4 nb0[K], pl0;
  {
    const int i0 = (tile_of(wid) << 5) + lane;
    const bool pre = i0 < n_ub;
#pragma unroll
    for (int j = 0; j < K; j++)
      nb0[j] = pre ? __ldg(&nbr[(size_t)i0 * K + j]) : make_float4(0.f, 0.f, 0.f, 0.f);
    pl0 = pre ? __ldg(&scan_L[i0]) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int n = st->n_scan;
  const int ntiles = (n + 31) >> 5;
  __syncthreads();
  tl_mark(tl, 2);
  // phase-2 role of this lane: group of points, block (bi, bj), bi <= bj
  constexpr int ROLES = NB * (NB + 1) / 2;
  constexpr int GROUPS = 32 / ROLES;
  const bool p2_lane = lane < ROLES * GROUPS;

  const int nrounds = (ntiles + nb - 1) / nb;
  for (int r = wid; r < nrounds; r += WARPS) {
    const int tile = tile_of(r);
    const int i = (tile << 5) + lane; // >= n for the tiles of the last round that do not exist: the warp idles through
    // ---- phase 1 ----
    if (r != wid && i < n) {
#pragma unroll
      for (int j = 0; j < K; j++)
        nb0[j] = __ldg(&nbr[(size_t)i * K + j]);
      pl0 = __ldg(&scan_L[i]);
    }
    int status = 1;
    double g[NG];
    RowOut ro;
    ro.W = ro.V = ro.chi2 = ro.rss = ro.raw_n = ro.res5 = ro.d = ro.sum_rp = 0.0;
    ro.n[0] = ro.n[1] = ro.n[2] = 0.0;
#pragma unroll
    for (int a = 0; a < NG; a++)
      g[a] = 0.0;
    // the kNN kernel marks a query get_neighbors rejects (LidarHelper.cpp:280-285) with a NaN first coordinate
    if (i < n && nb0[0].x == nb0[0].x)
      status = point_row<K, NB, EXPORT>(rc, pl0, nb0, g, ro);
    const bool okp = status == 0;
    if (EXPORT && i < n) { // parity export only: whitened g like the reference's H row
      PointRow pr;
      const bool reached = status == 0 || status == 4;
      const double ir = reached ? 1.0 / ro.raw_n : 0.0;
#pragma unroll
      for (int a = 0; a < ML_G; a++)
        pr.g[a] = (a < NG && reached) ? g[a < NG ? a : 0] * ir : 0.0;
      pr.w = reached ? ro.W * ro.raw_n * ro.raw_n : 0.0;
      pr.vr = reached ? ro.V * ro.raw_n : 0.0;
      pr.chi2 = reached ? ro.chi2 : 0.0;
      pr.res = reached ? ro.res5 * ir : 0.0;
      pr.raw_n = reached ? ro.raw_n : 0.0;
      pr.rss = reached ? ro.rss : 0.0;
      const bool has_plane = status == 0 || status == 3 || status == 4;
      pr.plane[0] = has_plane ? ro.n[0] : 0.0, pr.plane[1] = has_plane ? ro.n[1] : 0.0, pr.plane[2] = has_plane ? ro.n[2] : 0.0;
      pr.plane[3] = has_plane ? ro.d : 0.0;
      pr.status = status;
      pr.pad = 0;
      rows_out[i] = pr;
    }
#pragma unroll
    for (int a = 0; a < NG; a++)
      sg[lane * ROW + a] = okp ? g[a] : 0.0;
    sw[lane] = okp ? ro.W : 0.0;
    sv[lane] = okp ? ro.V : 0.0;
    {
      const bool tested = (status == 0 || status == 4) && (ro.chi2 == ro.chi2);
      double vchi = tested ? ro.chi2 : 0.0, vrss = tested ? ro.rss : 0.0, vsr = (EXPORT && tested) ? ro.sum_rp : 0.0;
      double vchi2 = vchi * vchi;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        vchi += __shfl_xor_sync(0xffffffffu, vchi, o);
        vchi2 += __shfl_xor_sync(0xffffffffu, vchi2, o);
        vrss += __shfl_xor_sync(0xffffffffu, vrss, o);
        if (EXPORT)
          vsr += __shfl_xor_sync(0xffffffffu, vsr, o);
      }
      if (lane == 0) {
        sstat[0] += vchi;
        sstat[1] += vchi2;
        sstat[2] += vrss;
        if (EXPORT)
          sstat[6] += vsr; // slots 3-5 hold the integer status counters
      }
      // status counters: lane q < 5 counts status q
      int mine = 0;
#pragma unroll
      for (int q = 0; q < 5; q++) {
        const int cq = __popc(__ballot_sync(0xffffffffu, i < n && status == q));
        mine = lane == q ? cq : mine;
      }
      if (lane < 5)
        scst[lane] += mine;
    }
    const unsigned any_ok = __ballot_sync(0xffffffffu, okp);
    __syncwarp();
    // ---- phase 2 ----
    if (any_ok && p2_lane) {
      const int grp = lane / ROLES, role = lane - grp * ROLES;
      int bi = 0, bj = 0;
      {
        int rr = 0, rem = role;
#pragma unroll
        for (int q = 0; q < NB; q++) {
          if (rem >= NB - rr) {
            rem -= NB - rr;
            rr++;
          }
        }
        bi = rr;
        bj = rr + rem;
      }
      double a9[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      double b3[3] = {0, 0, 0};
      const double *gI = sg + 3 * bi, *gJ = sg + 3 * bj;
#pragma unroll 4
      for (int p = grp; p < 32; p += GROUPS) {
        const double w = sw[p];
        const double x0 = gI[p * ROW] * w, x1 = gI[p * ROW + 1] * w, x2 = gI[p * ROW + 2] * w;
        const double y0 = gJ[p * ROW], y1 = gJ[p * ROW + 1], y2 = gJ[p * ROW + 2];
        a9[0] += x0 * y0, a9[1] += x0 * y1, a9[2] += x0 * y2;
        a9[3] += x1 * y0, a9[4] += x1 * y1, a9[5] += x1 * y2;
        a9[6] += x2 * y0, a9[7] += x2 * y1, a9[8] += x2 * y2;
        if (bi == bj) {
          const double v = sv[p];
          b3[0] += y0 * v, b3[1] += y1 * v, b3[2] += y2 * v;
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
  __syncwarp();
  tl_mark(tl, 3);
  if (lane < 5 && scst[lane])
    atomicAdd(&scnt[lane], scst[lane]);
  __syncthreads();
  tl_mark(tl, 4);
  // ---- CTA partial: sum over the warps in warp order. Entry e < NSYM: G(gi, gj), gi <= gj, by rows of the NG x NG upper
  //      triangle; NSYM <= e < NE: b; then the three statistics. Straight into this CTA's block of the pinned result area ----
  RRPart *mine = parts + blockIdx.x;
  if (t < NE) {
    int erole, q;
    if (t < NSYM) {
      int r = 0, rem = t;
      while (rem >= NG - r) {
        rem -= NG - r;
        r++;
      }
      const int gi = r, gj = r + rem;
      erole = rr_role_of<NB>(gi / 3, gj / 3);
      q = (gi % 3) * 3 + (gj % 3);
    } else {
      const int a = t - NSYM;
      erole = rr_role_of<NB>(a / 3, a / 3);
      q = 9 + a % 3;
    }
    double v = 0.0;
    for (int w = 0; w < WARPS; w++)
#pragma unroll
      for (int gq = 0; gq < GROUPS; gq++)
        v += slabs[(size_t)w * SLAB + 32 * ROW + 64 + (gq * ROLES + erole) * RR_RACC + q];
    mine->v[t] = v;
  } else if (t < NRED) {
    const int q = t - NE == 3 ? 6 : t - NE; // the fourth statistic (sum_res) sits behind the integer counters in the slab
    double v = 0.0;
    for (int w = 0; w < WARPS; w++)
      v += slabs[(size_t)w * SLAB + 32 * ROW + 64 + 32 * RR_RACC + q];
    mine->v[t] = v;
  } else if (t >= 192 && t < 197) {
    mine->c[t - 192] = scnt[t - 192];
  } else if (t == 197) {
    mine->n_scan = n;
  }
  tl_mark(tl, 5);
}

// parity export: h = T^T g for every point (n x hx)
__global__ void __launch_bounds__(256) k_expand_h(const ScanState *__restrict__ st, const PointRow *__restrict__ rows,
                                                 const double *__restrict__ T /* [ML_G][ML_MAX_HX] */, int hx,
                                                 double *__restrict__ h_rows) {
  const long long tot = (long long)st->n_scan * hx;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < tot; idx += (long long)gridDim.x * blockDim.x) {
    const int p = (int)(idx / hx), j = (int)(idx - (long long)p * hx);
    double s = 0.0;
    if (rows[p].status == 0 || rows[p].status == 4) {
#pragma unroll
      for (int l = 0; l < ML_G; l++)
        s += rows[p].g[l] * T[l * ML_MAX_HX + j];
    }
    h_rows[idx] = s;
  }
}

} // namespace ml
