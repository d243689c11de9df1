// ICP pre-alignment of the scan against the local map: the default libpointmatcher chain
// that LidarHelper::transform_to_map configures (mins/src/update/lidar/LidarHelper.cpp:156-187;
// thirdparty/libpointmatcher/pointmatcher/ICP.cpp:100-113 setDefault, :265-449 compute), on the device:
//   reference : map points with ||p|| < map_icp_dist (DataPointsFilters/MaxDist.cpp), kept with probability `ratio`, each
//               with a surface normal (DataPointsFilters/SamplingSurfaceNormal.cpp). The reference library bins the cloud by
//               recursive median splits until <= 7 points and takes the PCA normal of the bin; here the normal is the PCA
//               normal of the 7 nearest map points found on the kNN grid (same support size, no sequential recursion).
//   reading   : the scan cloud, kept with probability 0.75 (DataPointsFilters/RandomSampling.cpp:92-109).
//   iteration : rigid transform -> exact 1-NN (MatchersImpl.cpp:78-101, libnabo eps = 0) -> TrimmedDist 85 % quantile
//               (OutlierFiltersImpl.cpp:140-147, Matches.cpp:61-87) -> point-to-plane normal equations
//               (ErrorMinimizers/PointToPlane.cpp:172-260); the 6x6 solve and the convergence checks
//               (TransformationCheckersImpl.cpp: Counter 40, Differential 1e-3 / 1e-3 / smooth 3) run on the host.
// The reference's samplers are seeded from std::random_device / std::rand (non-deterministic); here every Bernoulli draw is
// Philox-4x32-10 keyed by the caller's seed with the point index as counter, so a run is reproducible.
#pragma once
#include "sim.cuh"
#include "update.cuh"
#include "rows.cuh"

namespace ml {

constexpr int ICP_NRED = 21 + 6 + 1; // A (upper triangle) | b | sum of squared point-to-plane errors
constexpr int ICP_MAX_PARTS = 1024;

struct IcpState {
  int n_ref;   // kept reference points that have a normal
  int n_read;  // reading points after RandomSampling
  int n_valid; // matches with a finite distance
  int n_used;  // matches with weight 1
  double mean[3]; // mean of the kept reference points (libpointmatcher's <refMean> frame origin)
  float limit;    // TrimmedDist threshold on the squared match distance
  uint32_t sel_prefix, sel_rank;
  uint32_t hist[256];
  unsigned int ticket, ticket2, ticket3;
  int pad;
};

// result block in pinned host memory, written by the last CTA of k_icp_normal_eq
struct IcpHostOut {
  double v[ICP_NRED];
  int n_ref, n_read, n_valid, n_used;
  float limit;
  int pad;
};

__device__ __forceinline__ float philox_u01(uint32_t idx, uint32_t stream, uint64_t seed) {
  uint32_t r[4];
  philox4x32_10(idx, stream, 0u, 0u, (uint32_t)(seed & 0xffffffffu), (uint32_t)(seed >> 32), r);
  return __fmul_rn((float)(r[0] >> 8), 5.9604644775390625e-08f); // [0, 1)
}

// ---- reference selection: MaxDist (Euclidean norm < maxDist, MaxDist.cpp) and the Bernoulli(ratio) draw of
// SamplingSurfaceNormal.cpp (samplingMethod 0) -----------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_icp_ref_select(const float4 *__restrict__ map_pts, const int *__restrict__ d_n, float max_dist,
                                                       float ratio, uint64_t seed, uint8_t *__restrict__ keep) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_n)
    return;
  const float4 p = map_pts[i];
  const bool in = norm3_ref(p.x, p.y, p.z) < max_dist;
  keep[i] = (in && philox_u01((uint32_t)i, 1u, seed) < ratio) ? 1 : 0;
}

// unit eigenvector of the smallest eigenvalue of the symmetric 3x3 covariance C (row-major upper: c00 c01 c02 c11 c12 c22).
// Returns false when the matrix has rank < 2 (SamplingSurfaceNormal.cpp drops such bins as "unfit").
__device__ __forceinline__ bool smallest_eigvec(double c00, double c01, double c02, double c11, double c12, double c22, double n[3]) {
  double lmin, lmax;
  sym3_eig_minmax(c00, c01, c02, c11, c12, c22, &lmin, &lmax);
  const double tr = c00 + c11 + c22;
  const double lmid = tr - lmin - lmax;
  if (!(lmax > 0.0) || !(lmid > 1e-12 * lmax))
    return false;
  // rows of (C - lmin I); the eigenvector is orthogonal to all of them: take the best-conditioned cross product
  const double r0[3] = {c00 - lmin, c01, c02}, r1[3] = {c01, c11 - lmin, c12}, r2[3] = {c02, c12, c22 - lmin};
  double x0[3], x1[3], x2[3];
  d_cross(r0, r1, x0);
  d_cross(r0, r2, x1);
  d_cross(r1, r2, x2);
  const double n0 = x0[0] * x0[0] + x0[1] * x0[1] + x0[2] * x0[2];
  const double n1 = x1[0] * x1[0] + x1[1] * x1[1] + x1[2] * x1[2];
  const double n2 = x2[0] * x2[0] + x2[1] * x2[1] + x2[2] * x2[2];
  const double *b = x0;
  double nb = n0;
  if (n1 > nb)
    b = x1, nb = n1;
  if (n2 > nb)
    b = x2, nb = n2;
  if (!(nb > 0.0))
    return false;
  const double inv = 1.0 / sqrt(nb);
  n[0] = b[0] * inv, n[1] = b[1] * inv, n[2] = b[2] * inv;
  return true;
}

// ---- normals of the kept reference points: exact NK-NN on the map's own grid, PCA of the neighbourhood -----------------
// Warps stride over groups of 32 grid cells; the points of a non-empty cell are processed 32 at a time as one kNN unit
// (they are contiguous in cell_pts). Points without NK neighbours inside `radius` or with a degenerate neighbourhood are
// dropped from the reference.
template <int NK>
__global__ void __launch_bounds__(KNN_WARPS * 32)
    k_icp_normals(const GridDesc *__restrict__ gd, const uint32_t *__restrict__ cell_start, const float4 *__restrict__ cell_pts,
                  const float4 *__restrict__ map_pts, float radius2, uint8_t *__restrict__ keep, float4 *__restrict__ normals) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const GridDesc g = *gd;
  const int R = max(1, (int)ceil(sqrt((double)radius2) * (1.0 + 1e-5) * g.inv_cell));
  const int ngroups = (g.ncells + 31) >> 5;
  for (int grp = warp; grp < ngroups; grp += nwarps) {
    const int c = (grp << 5) + lane;
    uint32_t beg = 0, end = 0;
    if (c < g.ncells) {
      beg = cell_start[c];
      end = cell_start[c + 1];
    }
    unsigned m = __ballot_sync(0xffffffffu, end > beg);
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1;
      const uint32_t cb = __shfl_sync(0xffffffffu, beg, src), ce = __shfl_sync(0xffffffffu, end, src);
      const int cc = (grp << 5) + src;
      KnnUnit un;
      un.cx = cc % g.dimx;
      un.cy = (cc / g.dimx) % g.dimy;
      un.cz = cc / (g.dimx * g.dimy);
      for (uint32_t base = cb; base < ce; base += 32) {
        un.qbase = (int)base;
        un.cnt = (int)min(32u, ce - base);
        const bool has = lane < un.cnt;
        const float4 q = cell_pts[base + (has ? lane : 0)];
        const int qi = __float_as_int(q.w);
        const bool active = has && keep[qi];
        if (!__any_sync(0xffffffffu, active))
          continue;
        TopK<NK> best;
        knn_walk<NK>(un, q, active, g, R, cell_start, cell_pts, best);
        if (active) {
          bool ok = best.i[NK - 1] != 0x7fffffff && best.d[NK - 1] <= radius2;
          double nrm[3] = {0, 0, 0};
          if (ok) {
            double P[NK][3], mx = 0, my = 0, mz = 0;
#pragma unroll
            for (int j = 0; j < NK; j++) {
              const float4 p = __ldg(&map_pts[best.i[j]]);
              P[j][0] = p.x, P[j][1] = p.y, P[j][2] = p.z;
              mx += P[j][0], my += P[j][1], mz += P[j][2];
            }
            mx /= NK, my /= NK, mz /= NK;
            double c00 = 0, c01 = 0, c02 = 0, c11 = 0, c12 = 0, c22 = 0;
#pragma unroll
            for (int j = 0; j < NK; j++) {
              const double x = P[j][0] - mx, y = P[j][1] - my, z = P[j][2] - mz;
              c00 += x * x, c01 += x * y, c02 += x * z, c11 += y * y, c12 += y * z, c22 += z * z;
            }
            ok = smallest_eigvec(c00 / NK, c01 / NK, c02 / NK, c11 / NK, c12 / NK, c22 / NK, nrm);
          }
          if (ok)
            normals[qi] = make_float4((float)nrm[0], (float)nrm[1], (float)nrm[2], 0.f);
          else
            keep[qi] = 0;
        }
      }
    }
  }
}

// ---- matcher structure: the map's cell-ordered point array with the unselected points moved to infinity (same
// cell_start table, no second grid build) -------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_icp_mask_cells(const float4 *__restrict__ cell_pts, const int *__restrict__ d_n,
                                                       const uint8_t *__restrict__ keep, float4 *__restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= *d_n)
    return;
  float4 p = cell_pts[j];
  if (!keep[__float_as_int(p.w)])
    p.x = p.y = p.z = INFINITY;
  out[j] = p;
}

// fixed-order block reduction of NV doubles per thread; result valid in thread 0
template <int NV, int BLOCK> __device__ __forceinline__ void block_sum(double (&v)[NV], double *smem /* [BLOCK/32][NV] */) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < NV; q++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
    if (lane == 0)
      smem[wid * NV + q] = v[q];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int q = 0; q < NV; q++) {
      double s = 0.0;
      for (int w = 0; w < BLOCK / 32; w++)
        s += smem[w * NV + q];
      v[q] = s;
    }
  }
  __syncthreads();
}

// mean and count of the kept reference points (ICP.cpp:292-300), deterministic: grid-stride partials, last CTA adds them
// in CTA order
__global__ void __launch_bounds__(256) k_icp_ref_mean(const float4 *__restrict__ map_pts, const int *__restrict__ d_n,
                                                     const uint8_t *__restrict__ keep, double *__restrict__ parts /* [grid][4] */,
                                                     IcpState *__restrict__ S) {
  __shared__ double sm[8 * 4];
  __shared__ unsigned int s_last;
  const int n = *d_n;
  double v[4] = {0, 0, 0, 0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (keep[i]) {
      const float4 p = map_pts[i];
      v[0] += (double)p.x, v[1] += (double)p.y, v[2] += (double)p.z, v[3] += 1.0;
    }
  }
  block_sum<4, 256>(v, sm);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int q = 0; q < 4; q++)
      parts[blockIdx.x * 4 + q] = v[q];
    __threadfence();
    s_last = (atomicAdd(&S->ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (!s_last || threadIdx.x != 0)
    return;
  __threadfence();
  const volatile double *vp = parts;
  double s[4] = {0, 0, 0, 0};
  for (unsigned b = 0; b < gridDim.x; b++)
    for (int q = 0; q < 4; q++)
      s[q] += vp[b * 4 + q];
  const double cnt = s[3] > 0.0 ? s[3] : 1.0;
  S->mean[0] = s[0] / cnt, S->mean[1] = s[1] / cnt, S->mean[2] = s[2] / cnt;
  S->n_ref = (int)s[3];
  S->ticket = 0u;
}

// ---- reading selection (RandomSampling.cpp:92-109): Bernoulli(prob) flags; the compaction keeps at most
// floor(n * prob) + 1 points like the reference's `j <= nbPointsOut` loop bound -----------------------------------------
__global__ void __launch_bounds__(256) k_icp_read_flags(const int *__restrict__ d_n, float prob, uint64_t seed, uint32_t *__restrict__ flag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_n)
    return;
  flag[i] = philox_u01((uint32_t)i, 2u, seed) < prob ? 1u : 0u;
}
__global__ void __launch_bounds__(256)
    k_icp_read_compact(const float4 *__restrict__ scan, const int *__restrict__ d_n, float prob, const uint32_t *__restrict__ flag,
                       const uint32_t *__restrict__ pos, float4 *__restrict__ out, IcpState *__restrict__ S, ScanState *__restrict__ ist) {
  const int n = *d_n;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t nb_out = (uint32_t)((float)n * prob);
  if (i == 0) {
    const uint32_t tot = n > 0 ? pos[n] : 0u;
    const int kept = (int)min(tot, nb_out + 1u);
    S->n_read = kept;
    ist->n_raw = ist->n_orig = ist->n_scan = kept; // the binning kernels read the query count from a ScanState
  }
  if (i >= n)
    return;
  if (flag[i] && pos[i] <= nb_out)
    out[pos[i]] = scan[i];
}

// reading -> map frame with the current estimate (TransformationsImpl.cpp:61-72), fp32
__global__ void __launch_bounds__(256) k_icp_transform(const float4 *__restrict__ in, const IcpState *__restrict__ S, Tf32 T,
                                                      float4 *__restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= S->n_read)
    return;
  out[i] = tf_apply(T, in[i]);
}

// ---- TrimmedDist: radix select of the floor(n_valid * ratio)-th smallest squared match distance, 8 bits per pass from
// the top byte (non-negative floats order like their bit patterns). One launch per pass; the last CTA of a pass narrows
// the prefix. pass = 3, 2, 1, 0. -----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_icp_select_pass(const float *__restrict__ d2, const uint8_t *__restrict__ ok, IcpState *__restrict__ S,
                                                        int pass, float ratio) {
  __shared__ uint32_t sh[256];
  __shared__ unsigned int s_last;
  sh[threadIdx.x] = 0;
  __syncthreads();
  const int n = S->n_read;
  const uint32_t prefix = S->sel_prefix;
  const int hi_shift = 8 * (pass + 1);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (!ok[i])
      continue;
    const uint32_t key = __float_as_uint(d2[i]);
    if (pass == 3 || (key >> hi_shift) == prefix)
      atomicAdd(&sh[(key >> (8 * pass)) & 255u], 1u);
  }
  __syncthreads();
  if (sh[threadIdx.x])
    atomicAdd(&S->hist[threadIdx.x], sh[threadIdx.x]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0)
    s_last = (atomicAdd(&S->ticket2, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (!s_last || threadIdx.x != 0)
    return;
  __threadfence();
  volatile uint32_t *hv = S->hist;
  uint32_t rank = S->sel_rank;
  if (pass == 3) {
    uint32_t tot = 0;
    for (int b = 0; b < 256; b++)
      tot += hv[b];
    S->n_valid = (int)tot;
    rank = ratio >= 1.0f ? (tot ? tot - 1u : 0u) : (uint32_t)((float)tot * ratio); // Matches.cpp:83-86
    if (tot && rank >= tot)
      rank = tot - 1u;
  }
  uint32_t cum = 0;
  int bin = 255;
  for (int b = 0; b < 256; b++) {
    const uint32_t hb = hv[b];
    if (rank < cum + hb) {
      bin = b;
      break;
    }
    cum += hb;
  }
  const uint32_t np = (pass == 3 ? 0u : (prefix << 8)) | (uint32_t)bin;
  S->sel_prefix = np;
  S->sel_rank = rank - cum;
  if (pass == 0)
    S->limit = S->n_valid > 0 ? __uint_as_float(np) : -1.0f;
  for (int b = 0; b < 256; b++)
    hv[b] = 0u;
  S->ticket2 = 0u;
}

// ---- point-to-plane normal equations (PointToPlane.cpp:172-247): F = [p x n ; n], A = sum w F F^T, b = -sum w F ((p - r).n),
// in the <refMean> frame (coordinates relative to the reference mean), fp64 accumulation, fixed order --------------------
__global__ void __launch_bounds__(256)
    k_icp_normal_eq(const float4 *__restrict__ q_map, const int32_t *__restrict__ nn_idx, const float *__restrict__ nn_d2,
                    const uint8_t *__restrict__ nn_ok, const float4 *__restrict__ map_pts, const float4 *__restrict__ normals,
                    IcpState *__restrict__ S, double *__restrict__ parts /* [grid][ICP_NRED + 1] */, IcpHostOut *__restrict__ out) {
  __shared__ double sm[8 * (ICP_NRED + 1)];
  __shared__ unsigned int s_last;
  const int n = S->n_read;
  const float limit = S->limit;
  const double mx = S->mean[0], my = S->mean[1], mz = S->mean[2];
  double v[ICP_NRED + 1];
#pragma unroll
  for (int q = 0; q <= ICP_NRED; q++)
    v[q] = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (!nn_ok[i] || !(nn_d2[i] <= limit))
      continue;
    const int j = nn_idx[i];
    const float4 qp = q_map[i], rp = map_pts[j], nf = normals[j];
    const double p[3] = {(double)qp.x - mx, (double)qp.y - my, (double)qp.z - mz};
    const double nn[3] = {nf.x, nf.y, nf.z};
    double F[6];
    d_cross(p, nn, F);
    F[3] = nn[0], F[4] = nn[1], F[5] = nn[2];
    const double dot = ((double)qp.x - (double)rp.x) * nn[0] + ((double)qp.y - (double)rp.y) * nn[1] + ((double)qp.z - (double)rp.z) * nn[2];
    int e = 0;
#pragma unroll
    for (int a = 0; a < 6; a++)
#pragma unroll
      for (int b = a; b < 6; b++)
        v[e++] += F[a] * F[b];
#pragma unroll
    for (int a = 0; a < 6; a++)
      v[21 + a] -= F[a] * dot;
    v[27] += dot * dot;
    v[28] += 1.0;
  }
  block_sum<ICP_NRED + 1, 256>(v, sm);
  if (threadIdx.x == 0) {
    for (int q = 0; q <= ICP_NRED; q++)
      parts[blockIdx.x * (ICP_NRED + 1) + q] = v[q];
    __threadfence();
    s_last = (atomicAdd(&S->ticket3, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (!s_last)
    return;
  __threadfence();
  const volatile double *vp = parts;
  if (threadIdx.x <= ICP_NRED) {
    double s = 0.0;
    for (unsigned b = 0; b < gridDim.x; b++)
      s += vp[b * (ICP_NRED + 1) + threadIdx.x];
    if (threadIdx.x < ICP_NRED)
      out->v[threadIdx.x] = s;
    else {
      S->n_used = (int)s;
      out->n_used = (int)s;
      out->n_ref = S->n_ref, out->n_read = S->n_read, out->n_valid = S->n_valid;
      out->limit = limit;
      S->ticket3 = 0u;
      S->sel_prefix = 0u, S->sel_rank = 0u;
    }
  }
  __threadfence_system();
}

} // namespace ml
