// Simulator LiDAR model on the device: rays x planes nearest-hit with the reference's in-quad test,
// range gate, fp32 noise from a counter-based generator, stable compaction in ray order.
// Reference: mins/src/sim/Simulator.cpp:742-815 (get_lidar_pointcloud), mins/src/sim/SimulationPlane.h:43-95
// (plane from three corners, calculate_intersection), Simulator.cpp:990-1096 (planes from the floor plan; built on the
// host, mins_b200/sim/world.py).
#pragma once
#include "scan.cuh"

namespace ml {

// one plane = 22 doubles: [A B C D | top_left | bottom_right | V1 V2 V3 V4] (V* = normalised edge directions,
// SimulationPlane.h:80-87). Only the SIGN of U.V is used (:94), so U is not normalised here: U/|U| . V > 0 <=> U . V > 0,
// and |U| = 0 gives NaN > 0 = false there and 0 > 0 = false here.
constexpr int SIM_PLANE_STRIDE = 22;
constexpr int SIM_BLOCK = 256;
constexpr int SIM_PLANES_SMEM = 256; // planes staged per pass (44 KB of doubles + 2 KB numerators)

struct SimPose {
  double o[3];  // LiDAR origin in the world: p_IinG + R_GtoI^T p_LinI (Simulator.cpp:765)
  double R[9];  // R_LtoG = R_GtoI^T R_ItoL^T (:769), row-major
};

// Philox-4x32-10 (Salmon et al. 2011): counter = (ray index, lane, 0, 0), key = seed. Restated in
// oracle/pyoracle.py::philox4x32 for the parity test.
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
  for (int r = 0; r < 10; r++) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0, c1 = n1, c2 = n2, c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0, out[1] = c1, out[2] = c2, out[3] = c3;
}
// Box-Muller in fp32: u in (0, 1] from the top 24 bits
__device__ __forceinline__ float u01_open(uint32_t x) { return __fmul_rn((float)((x >> 8) + 1u), 5.9604644775390625e-08f); }
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float *z0, float *z1) {
  const float r = sqrtf(__fmul_rn(-2.0f, logf(u01_open(a))));
  const float th = __fmul_rn(6.283185307179586f, u01_open(b));
  *z0 = __fmul_rn(r, cosf(th));
  *z1 = __fmul_rn(r, sinf(th));
}

// pass 1: nearest hit distance per ray + per-tile count of rays inside [min_range, max_range]
// SWEEP: motion distortion of Simulator.cpp:761-771 (raw_remove_motion_blur): ray i of `total` is cast from the pose at
// time - (total - i) * raw_point_dt. The pose track arrives as n_knots poses [R_LtoG (9) | p_LinG (3)] sampled uniformly over
// [time - total * raw_point_dt, time]; a ray blends its two bracketing knots linearly (the same rule as k_deskew).
template <bool SWEEP>
__global__ void __launch_bounds__(SIM_BLOCK)
    k_ray_cast(const double *__restrict__ planes, int n_planes, const double *__restrict__ dirs_L, int n_rays, SimPose pose,
               const double *__restrict__ knots, int n_knots, double rmin, double rmax, double *__restrict__ dist,
               uint32_t *__restrict__ tile_sums) {
  __shared__ double sp[SIM_PLANES_SMEM * SIM_PLANE_STRIDE];
  __shared__ double snum[SIM_PLANES_SMEM];
  __shared__ uint32_t wsum[SIM_BLOCK / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  double best[RF_TILE / SIM_BLOCK];
  double dg[RF_TILE / SIM_BLOCK][3];
  double og[SWEEP ? RF_TILE / SIM_BLOCK : 1][3]; // per-ray origin (SWEEP only)
  const int base = blockIdx.x * RF_TILE;
#pragma unroll
  for (int r = 0; r < RF_TILE / SIM_BLOCK; r++) {
    const int i = base + r * SIM_BLOCK + threadIdx.x;
    best[r] = INFINITY;
    double dx = 0, dy = 0, dz = 0;
    if (i < n_rays)
      dx = dirs_L[3 * (size_t)i], dy = dirs_L[3 * (size_t)i + 1], dz = dirs_L[3 * (size_t)i + 2];
    if (SWEEP) {
      double P[12];
      const double u = (double)min(i, n_rays - 1) / (double)n_rays * (double)(n_knots - 1);
      const int k0 = min((int)u, n_knots - 2);
      const double a = u - (double)k0;
#pragma unroll
      for (int e = 0; e < 12; e++)
        P[e] = (1.0 - a) * knots[(size_t)k0 * 12 + e] + a * knots[(size_t)(k0 + 1) * 12 + e];
      dg[r][0] = P[0] * dx + P[1] * dy + P[2] * dz;
      dg[r][1] = P[3] * dx + P[4] * dy + P[5] * dz;
      dg[r][2] = P[6] * dx + P[7] * dy + P[8] * dz;
      og[r][0] = P[9], og[r][1] = P[10], og[r][2] = P[11];
    } else {
      dg[r][0] = pose.R[0] * dx + pose.R[1] * dy + pose.R[2] * dz;
      dg[r][1] = pose.R[3] * dx + pose.R[4] * dy + pose.R[5] * dz;
      dg[r][2] = pose.R[6] * dx + pose.R[7] * dy + pose.R[8] * dz;
    }
  }
  for (int p0 = 0; p0 < n_planes; p0 += SIM_PLANES_SMEM) {
    const int np = min(SIM_PLANES_SMEM, n_planes - p0);
    __syncthreads();
    for (int e = threadIdx.x; e < np * SIM_PLANE_STRIDE; e += SIM_BLOCK)
      sp[e] = planes[(size_t)p0 * SIM_PLANE_STRIDE + e];
    __syncthreads();
    for (int p = threadIdx.x; p < np; p += SIM_BLOCK) {
      const double *q = sp + p * SIM_PLANE_STRIDE;
      snum[p] = -(q[0] * pose.o[0] + q[1] * pose.o[1] + q[2] * pose.o[2] + q[3]); // SimulationPlane.h:76 numerator
    }
    __syncthreads();
    for (int p = 0; p < np; p++) {
      const double *q = sp + p * SIM_PLANE_STRIDE;
      const double A = q[0], B = q[1], C = q[2];
#pragma unroll
      for (int r = 0; r < RF_TILE / SIM_BLOCK; r++) {
        const double ox = SWEEP ? og[r][0] : pose.o[0], oy = SWEEP ? og[r][1] : pose.o[1], oz = SWEEP ? og[r][2] : pose.o[2];
        const double num = SWEEP ? -(A * ox + B * oy + C * oz + q[3]) : snum[p];
        const double t = num / (A * dg[r][0] + B * dg[r][1] + C * dg[r][2]);
        const double px = ox + t * dg[r][0], py = oy + t * dg[r][1], pz = oz + t * dg[r][2];
        const double u1x = px - q[4], u1y = py - q[5], u1z = pz - q[6];
        const double u2x = px - q[7], u2y = py - q[8], u2z = pz - q[9];
        const bool hit = t > 0.0 && (u1x * q[10] + u1y * q[11] + u1z * q[12]) > 0.0 && (u1x * q[13] + u1y * q[14] + u1z * q[15]) > 0.0 &&
                         (u2x * q[16] + u2y * q[17] + u2z * q[18]) > 0.0 && (u2x * q[19] + u2y * q[20] + u2z * q[21]) > 0.0;
        if (hit)
          best[r] = fmin(best[r], t);
      }
    }
  }
  uint32_t cnt = 0;
#pragma unroll
  for (int r = 0; r < RF_TILE / SIM_BLOCK; r++) {
    const int i = base + r * SIM_BLOCK + threadIdx.x;
    const bool keep = i < n_rays && !(best[r] > rmax || best[r] < rmin); // Simulator.cpp:788
    if (i < n_rays)
      dist[i] = keep ? best[r] : -1.0;
    cnt += __popc(__ballot_sync(0xffffffffu, keep));
  }
  if (lane == 0)
    wsum[wid] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t s = 0;
#pragma unroll
    for (int w = 0; w < SIM_BLOCK / 32; w++)
      s += wsum[w];
    tile_sums[blockIdx.x] = s;
  }
}

// pass 2: stable compaction in ray order; p = d * dir_L cast to fp32 (+ raw_noise * N(0,1) per axis, Simulator.cpp:792-804)
__global__ void __launch_bounds__(SIM_BLOCK)
    k_ray_emit(const double *__restrict__ dist, const double *__restrict__ dirs_L, int n_rays, const uint32_t *__restrict__ tile_sums,
               float noise, uint32_t seed_lo, uint32_t seed_hi, float4 *__restrict__ out, int *__restrict__ d_n_out) {
  __shared__ uint32_t tile_prefix;
  __shared__ uint32_t rsum[RF_TILE / SIM_BLOCK][SIM_BLOCK / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (wid == 0) {
    uint32_t s = 0;
    for (int b = lane; b < (int)blockIdx.x; b += 32)
      s += tile_sums[b];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0)
      tile_prefix = s;
  }
  const int base = blockIdx.x * RF_TILE;
  double d[RF_TILE / SIM_BLOCK];
  uint32_t bal[RF_TILE / SIM_BLOCK];
#pragma unroll
  for (int r = 0; r < RF_TILE / SIM_BLOCK; r++) {
    const int i = base + r * SIM_BLOCK + threadIdx.x;
    d[r] = i < n_rays ? dist[i] : -1.0;
    bal[r] = __ballot_sync(0xffffffffu, d[r] >= 0.0);
    if (lane == 0)
      rsum[r][wid] = __popc(bal[r]);
  }
  __syncthreads();
  // order inside the tile: round r, then warp, then lane == ray index order
  uint32_t pos = tile_prefix;
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < RF_TILE / SIM_BLOCK; r++) {
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < SIM_BLOCK / 32; w++) {
      const uint32_t c = rsum[r][w];
      before += w < wid ? c : 0u;
      total += c;
    }
    if ((bal[r] >> lane) & 1u) {
      const int i = base + r * SIM_BLOCK + threadIdx.x;
      const uint32_t o = pos + before + __popc(bal[r] & lt);
      float x = (float)(d[r] * dirs_L[3 * (size_t)i]), y = (float)(d[r] * dirs_L[3 * (size_t)i + 1]),
            z = (float)(d[r] * dirs_L[3 * (size_t)i + 2]);
      if (noise > 0.f) {
        uint32_t rn[4];
        philox4x32_10((uint32_t)i, 0u, 0u, 0u, seed_lo, seed_hi, rn);
        float z0, z1, z2, z3;
        box_muller(rn[0], rn[1], &z0, &z1);
        box_muller(rn[2], rn[3], &z2, &z3);
        x = __fadd_rn(x, __fmul_rn(noise, z0));
        y = __fadd_rn(y, __fmul_rn(noise, z1));
        z = __fadd_rn(z, __fmul_rn(noise, z2));
      }
      out[o] = make_float4(x, y, z, 0.f);
    }
    pos += total;
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0)
    *d_n_out = (int)pos;
}

} // namespace ml
