// Host orchestration of ml_scan_icp (included by api.cu at namespace scope). Device work: icp.cuh. The per-iteration host
// part is what libpointmatcher does on 6x6 / 4x4 matrices: Cholesky solve (PointToPlane.cpp:140-143), angle-axis update
// (:258-283), T_iter composition (ICP.cpp:404) and the two transformation checkers (TransformationCheckersImpl.cpp).

namespace {

struct M4 {
  double m[16];
};
inline M4 m4_identity() {
  M4 r;
  for (int i = 0; i < 16; i++)
    r.m[i] = (i % 5 == 0) ? 1.0 : 0.0;
  return r;
}
inline M4 m4_mul(const M4 &a, const M4 &b) {
  M4 r;
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      double s = 0.0;
      for (int k = 0; k < 4; k++)
        s += a.m[4 * i + k] * b.m[4 * k + j];
      r.m[4 * i + j] = s;
    }
  return r;
}
inline M4 m4_rigid_inverse(const M4 &a) {
  M4 r = m4_identity();
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      r.m[4 * i + j] = a.m[4 * j + i];
  for (int i = 0; i < 3; i++)
    r.m[4 * i + 3] = -(r.m[4 * i] * a.m[3] + r.m[4 * i + 1] * a.m[7] + r.m[4 * i + 2] * a.m[11]);
  return r;
}
// Cholesky solve of the symmetric 6x6 system; false when A is not positive definite
inline bool chol6_solve(const double A[36], const double b[6], double x[6]) {
  double L[36] = {0};
  for (int i = 0; i < 6; i++)
    for (int j = 0; j <= i; j++) {
      double s = A[6 * i + j];
      for (int k = 0; k < j; k++)
        s -= L[6 * i + k] * L[6 * j + k];
      if (i == j) {
        if (!(s > 0.0))
          return false;
        L[6 * i + i] = std::sqrt(s);
      } else {
        L[6 * i + j] = s / L[6 * j + j];
      }
    }
  double y[6];
  for (int i = 0; i < 6; i++) {
    double s = b[i];
    for (int k = 0; k < i; k++)
      s -= L[6 * i + k] * y[k];
    y[i] = s / L[6 * i + i];
  }
  for (int i = 5; i >= 0; i--) {
    double s = y[i];
    for (int k = i + 1; k < 6; k++)
      s -= L[6 * k + i] * x[k];
    x[i] = s / L[6 * i + i];
  }
  return true;
}
inline double rot_angle_between(const M4 &a, const M4 &b) { // angle of R_a R_b^T (Quaternion::angularDistance)
  double tr = 0.0;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      tr += a.m[4 * i + j] * b.m[4 * i + j];
  double c = (tr - 1.0) * 0.5;
  c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
  return std::acos(c);
}

} // namespace

static int icp_alloc(ml_handle *h) {
  if (h->icp_keep)
    return ML_OK;
  const size_t ca = (size_t)h->cap_all, cs = (size_t)h->cap_scan;
  CK(cudaMalloc(&h->icp_keep, ca));
  CK(cudaMalloc(&h->icp_normals, sizeof(float4) * ca));
  CK(cudaMalloc(&h->icp_cell_pts, sizeof(float4) * ca));
  CK(cudaMalloc(&h->icp_read, sizeof(float4) * cs));
  CK(cudaMalloc(&h->icp_q, sizeof(float4) * cs));
  CK(cudaMalloc(&h->d_icp, sizeof(IcpState)));
  CK(cudaMalloc(&h->d_ist, sizeof(ScanState)));
  CK(cudaMemset(h->d_ist, 0, sizeof(ScanState)));
  CK(cudaMalloc(&h->icp_parts, sizeof(double) * ICP_MAX_PARTS * (ICP_NRED + 1)));
  CK(cudaMallocHost(&h->h_icp, sizeof(IcpHostOut)));
  return ML_OK;
}

extern "C" {

void ml_icp_default_options(ml_icp_options *o) {
  if (!o)
    return;
  o->max_dist = 20.0;       // OptionsLidar.h:97 map_icp_dist
  o->read_prob = 0.75;      // RandomSamplingDataPointsFilter prob
  o->ref_ratio = 0.5;       // SamplingSurfaceNormalDataPointsFilter ratio
  o->normal_knn = 7;        // SamplingSurfaceNormalDataPointsFilter knn
  o->trim_ratio = 0.85;     // TrimmedDistOutlierFilter ratio
  o->max_iter = 40;         // CounterTransformationChecker maxIterationCount
  o->min_diff_rot = 0.001;  // DifferentialTransformationChecker minDiffRotErr
  o->min_diff_trans = 0.001;
  o->smooth_len = 3;
  o->max_match_dist = 2.0;  // matcher bound (the reference's KDTreeMatcher is unbounded; see mins_lidar.h)
  o->normal_radius = 2.0;
  o->seed = 0;
}

int ml_scan_icp(ml_handle *h, const float T_init[16], const ml_icp_options *o, ml_icp_result *res) {
  if (!h || !T_init || !o || !res)
    return fail(h, ML_ERR_ARG, "ml_scan_icp: bad argument");
  RESOLVE(h);
  if (o->normal_knn != 7)
    return fail(h, ML_ERR_ARG, "ml_scan_icp: normal_knn must be 7 (the library default; the only instantiated kernel)");
  if (o->max_iter < 1 || o->smooth_len < 1 || !(o->max_match_dist > 0.0) || !(o->normal_radius > 0.0) || !(o->max_dist > 0.0))
    return fail(h, ML_ERR_ARG, "ml_scan_icp: bad options");
  if (!h->map_init)
    return fail(h, ML_SOFT_NO_MAP, "ml_scan_icp: map not initialised");
  int rc = icp_alloc(h);
  if (rc)
    return rc;
  h->rows_valid = false; // the matcher runs k_knn<1> into the kNN buffers the rows kernel read
  drop_stale_bins(h);
  memset(res, 0, sizeof(*res));
  for (int i = 0; i < 16; i++)
    res->T_LtoM[i] = res->T_icp[i] = T_init[i];
  const int M = h->map_n, n_ub = h->n_raw;
  if (n_ub <= 0 || M <= 0)
    return ML_SOFT_NO_ROWS;
  CK(cudaMemsetAsync(h->d_icp, 0, sizeof(IcpState), h->stream));
  // ---- reference: select, normals, matcher structure, mean ----
  const float nr2 = (float)(o->normal_radius * o->normal_radius), mm2 = (float)(o->max_match_dist * o->max_match_dist);
  LAUNCH(k_icp_ref_select, cdiv(M, 256), 256, h->map_pts[h->cur], h->d_map_n, (float)o->max_dist, (float)o->ref_ratio, o->seed, h->icp_keep);
  LAUNCH(k_icp_normals<7>, h->n_sm * 8, KNN_WARPS * 32, h->d_grid, h->cell_start, h->cell_pts, h->map_pts[h->cur], nr2, h->icp_keep,
         h->icp_normals);
  LAUNCH(k_icp_mask_cells, cdiv(M, 256), 256, h->cell_pts, h->d_map_n, h->icp_keep, h->icp_cell_pts);
  const int G = std::min(h->n_sm * 2, ICP_MAX_PARTS);
  LAUNCH(k_icp_ref_mean, G, 256, h->map_pts[h->cur], h->d_map_n, h->icp_keep, h->icp_parts, h->d_icp);
  // ---- reading: Bernoulli flags, stable compaction ----
  LAUNCH(k_icp_read_flags, cdiv(n_ub, 256), 256, ST(n_scan), (float)o->read_prob, o->seed, h->keep);
  excl_scan(h, h->keep, ST(n_scan), n_ub, h->pos, nullptr);
  LAUNCH(k_icp_read_compact, cdiv(n_ub, 256), 256, h->scan_pts, ST(n_scan), (float)o->read_prob, h->keep, h->pos, h->icp_read, h->d_icp,
         h->d_ist);
  IcpState S;
  CK(cudaMemcpyAsync(&S, h->d_icp, sizeof(IcpState), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  CK(cudaGetLastError());
  res->n_reference = S.n_ref;
  res->n_reading = S.n_read;
  if (S.n_ref < 6 || S.n_read < 6)
    return ML_SOFT_NO_ROWS; // libpointmatcher throws ConvergenceError; transform_to_map would not survive it either
  // ---- frames (ICP.cpp:292-300,351-353) ----
  M4 T_in = m4_identity();
  for (int i = 0; i < 16; i++)
    T_in.m[i] = (double)T_init[i];
  M4 T_refIn_refMean = m4_identity();
  T_refIn_refMean.m[3] = S.mean[0], T_refIn_refMean.m[7] = S.mean[1], T_refIn_refMean.m[11] = S.mean[2];
  const M4 T_refMean_dataIn = m4_mul(m4_rigid_inverse(T_refIn_refMean), T_in);
  M4 T_iter = m4_identity();
  std::vector<M4> hist;
  hist.push_back(T_iter);
  const uint32_t mask = (uint32_t)h->q_slots - 1u;
  const int n_read = S.n_read;
  const int Gs = std::max(1, std::min(h->n_sm, cdiv(n_read, 256)));
  bool iterate = true, max_reached = false, failed = false;
  int iters = 0;
  double last_err = 0.0;
  while (iterate) {
    const M4 Tc = m4_mul(T_refIn_refMean, m4_mul(T_iter, T_refMean_dataIn));
    Tf32 Tf;
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 4; c++)
        Tf.m[4 * r + c] = (float)Tc.m[4 * r + c];
    LAUNCH(k_icp_transform, cdiv(n_read, 256), 256, h->icp_read, h->d_icp, Tf, h->icp_q);
    const int ph = h->icp_knn_calls & 1;
    h->icp_knn_calls++;
    LAUNCH(k_query_bin, cdiv(n_read, 256), 256, h->icp_q, h->d_ist, h->d_grid, h->qtab, mask, h->q_slot, h->q_rank);
    LAUNCH(k_unit_alloc, cdiv(h->q_slots, 256), 256, h->qtab, h->q_slots, h->d_ist, ph, h->units, h->units_cap);
    LAUNCH(k_query_scatter, cdiv(n_read, 256), 256, h->icp_q, h->d_ist, h->qtab, h->q_slot, h->q_rank, h->q_sorted);
    LAUNCH(k_knn<1>, h->n_sm * 8, KNN_WARPS * 32, h->units, h->d_ist, ph, h->q_sorted, h->d_grid, h->cell_start, h->icp_cell_pts, mm2,
           h->knn_idx, h->knn_d2, h->knn_ok, h->qtab, h->q_slots, (const float4 *)nullptr, (float4 *)nullptr,
           (unsigned long long *)nullptr);
    for (int pass = 3; pass >= 0; pass--)
      LAUNCH(k_icp_select_pass, Gs, 256, h->knn_d2, h->knn_ok, h->d_icp, pass, (float)o->trim_ratio);
    LAUNCH(k_icp_normal_eq, Gs, 256, h->icp_q, h->knn_idx, h->knn_d2, h->knn_ok, h->map_pts[h->cur], h->icp_normals, h->d_icp, h->icp_parts,
           h->h_icp);
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    const IcpHostOut &R = *h->h_icp;
    res->n_matched = R.n_valid;
    res->n_used = R.n_used;
    if (R.n_used < 6) {
      failed = true;
      break;
    }
    double A[36], b[6], x[6];
    int e = 0;
    for (int a = 0; a < 6; a++)
      for (int c = a; c < 6; c++) {
        A[6 * a + c] = A[6 * c + a] = R.v[e];
        e++;
      }
    for (int a = 0; a < 6; a++)
      b[a] = R.v[21 + a];
    last_err = R.v[27] / (double)R.n_used;
    if (!chol6_solve(A, b, x)) {
      failed = true; // rank-deficient geometry: the reference falls back to a minimum-norm solution; here the prior is kept
      break;
    }
    // PointToPlane.cpp:262-283: rotation = AngleAxis(|x[0:3]|, x[0:3]/|..|), translation = x[3:6]
    M4 Td = m4_identity();
    const double ang = std::sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
    if (ang > 0.0) {
      const double ux = x[0] / ang, uy = x[1] / ang, uz = x[2] / ang, c = std::cos(ang), s = std::sin(ang), t = 1.0 - c;
      Td.m[0] = c + ux * ux * t, Td.m[1] = ux * uy * t - uz * s, Td.m[2] = ux * uz * t + uy * s;
      Td.m[4] = uy * ux * t + uz * s, Td.m[5] = c + uy * uy * t, Td.m[6] = uy * uz * t - ux * s;
      Td.m[8] = uz * ux * t - uy * s, Td.m[9] = uz * uy * t + ux * s, Td.m[10] = c + uz * uz * t;
    }
    Td.m[3] = x[3], Td.m[7] = x[4], Td.m[11] = x[5];
    T_iter = m4_mul(Td, T_iter);
    iters++;
    // CounterTransformationChecker, then DifferentialTransformationChecker
    if (iters >= o->max_iter) {
      iterate = false;
      max_reached = true;
      break;
    }
    hist.push_back(T_iter);
    if ((int)hist.size() > o->smooth_len) {
      double dr = 0.0, dt = 0.0;
      for (size_t i = hist.size() - 1; i >= hist.size() - (size_t)o->smooth_len; i--) {
        dr += std::fabs(rot_angle_between(hist[i], hist[i - 1]));
        const double ex = hist[i].m[3] - hist[i - 1].m[3], ey = hist[i].m[7] - hist[i - 1].m[7], ez = hist[i].m[11] - hist[i - 1].m[11];
        dt += std::sqrt(ex * ex + ey * ey + ez * ez);
      }
      dr /= o->smooth_len, dt /= o->smooth_len;
      if (dr < o->min_diff_rot && dt < o->min_diff_trans)
        iterate = false;
    }
  }
  const M4 T_icp = m4_mul(T_refIn_refMean, m4_mul(T_iter, T_refMean_dataIn));
  const M4 T_diff = m4_mul(T_icp, m4_rigid_inverse(T_in)); // LidarHelper.cpp:176
  const double shift = std::sqrt(T_diff.m[3] * T_diff.m[3] + T_diff.m[7] * T_diff.m[7] + T_diff.m[11] * T_diff.m[11]);
  res->iterations = iters;
  res->max_iter_reached = max_reached ? 1 : 0;
  res->mean_sq_err = last_err;
  for (int i = 0; i < 16; i++)
    res->T_icp[i] = (float)T_icp.m[i];
  res->success = (!failed && !max_reached && !(shift > 1.0)) ? 1 : 0; // LidarHelper.cpp:177-183
  if (res->success)
    for (int i = 0; i < 16; i++)
      res->T_LtoM[i] = res->T_icp[i];
  return ML_OK;
}

} // extern "C"
