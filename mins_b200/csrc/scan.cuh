// Scan pre-processing kernels: LiDARData construction (range filter), deskew, PCL-VoxelGrid down-sampling,
// rigid transform into the map frame. Reference citations relative to /root/reference/mins/src/update/lidar/.
#pragma once
#include "prims.cuh"

namespace ml {

// ---- LiDARData::LiDARData, LidarTypes.cpp:36-55 ------------------------------------------------------
// Stable compaction of the in-range points in two launches (no intermediate flag/prefix arrays in HBM):
//   k_range_count   : per-tile (2048 points) count of kept points; CTA 0 initialises the scan state
//   k_range_compact : every CTA sums the counts of the tiles before it (<= 128 values), ranks its own points with warp
//                     ballots and writes pointcloud, pointcloud_original and - when the map transform is already known -
//                     pointcloud_in_map (LidarHelper.cpp:192) in the same pass.
// Keep rule: norm in fp32 compared in double like `n > max_range || n < min_range` (LidarTypes.cpp:39).
constexpr int RF_BLOCK = 256;
constexpr int RF_TILE = 2048; // 8 warps x 8 rounds x 32 lanes

template <bool VEC4> __device__ __forceinline__ float4 load_raw(const float *__restrict__ raw, int stride, int i) {
  if (VEC4)
    return __ldg(reinterpret_cast<const float4 *>(raw) + i);
  const float *p = raw + (size_t)i * stride;
  return make_float4(p[0], p[1], p[2], 0.f);
}
__device__ __forceinline__ bool range_keep(const float4 p, double rmin, double rmax) {
  const float nrm = norm3_ref(p.x, p.y, p.z);
  return !((double)nrm > rmax || (double)nrm < rmin);
}

template <bool VEC4>
__global__ void __launch_bounds__(RF_BLOCK) k_range_count(const float *__restrict__ raw, int stride, int n, double rmin, double rmax,
                                                          ScanState *__restrict__ st, uint32_t *__restrict__ tile_sums) {
  __shared__ uint32_t wsum[RF_BLOCK / 32];
  pdl_enter();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int wbase = blockIdx.x * RF_TILE + wid * (RF_TILE / (RF_BLOCK / 32));
  uint32_t cnt = 0;
#pragma unroll
  for (int r = 0; r < 8; r++) {
    const int i = wbase + r * 32 + lane;
    const bool k = i < n && range_keep(load_raw<VEC4>(raw, stride, i), rmin, rmax);
    cnt += __popc(__ballot_sync(0xffffffffu, k));
  }
  if (lane == 0)
    wsum[wid] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t s = 0;
#pragma unroll
    for (int w = 0; w < RF_BLOCK / 32; w++)
      s += wsum[w];
    tile_sums[blockIdx.x] = s;
    if (blockIdx.x == 0) {
      st->n_raw = n;
      st->vox_n = 0;
      st->vox_passthrough = 0;
    }
  }
}

template <bool VEC4>
__global__ void __launch_bounds__(RF_BLOCK)
    k_range_compact(const float *__restrict__ raw, int stride, int n, float t_rel, double rmin, double rmax,
                    const uint32_t *__restrict__ tile_sums, ScanState *__restrict__ st, float4 *__restrict__ cloud,
                    float4 *__restrict__ orig, Tf32 T, float4 *__restrict__ in_map /* nullptr: transform later */) {
  __shared__ uint32_t wsum[RF_BLOCK / 32];
  __shared__ uint32_t tile_prefix;
  pdl_enter();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (wid == 0) { // counts of the tiles before this one, fixed order
    uint32_t s = 0;
    for (int b = lane; b < (int)blockIdx.x; b += 32)
      s += tile_sums[b];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0)
      tile_prefix = s;
  }
  const int wbase = blockIdx.x * RF_TILE + wid * (RF_TILE / (RF_BLOCK / 32));
  float4 pt[8];
  uint32_t bal[8];
  uint32_t cnt = 0;
#pragma unroll
  for (int r = 0; r < 8; r++) {
    const int i = wbase + r * 32 + lane;
    pt[r] = i < n ? load_raw<VEC4>(raw, stride, i) : make_float4(0.f, 0.f, 0.f, 0.f);
    const bool k = i < n && range_keep(pt[r], rmin, rmax);
    bal[r] = __ballot_sync(0xffffffffu, k);
    cnt += __popc(bal[r]);
  }
  if (lane == 0)
    wsum[wid] = cnt;
  __syncthreads();
  uint32_t pos = tile_prefix;
  for (int w = 0; w < wid; w++)
    pos += wsum[w];
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
    uint32_t tot = tile_prefix;
#pragma unroll
    for (int w = 0; w < RF_BLOCK / 32; w++)
      tot += wsum[w];
    st->n_orig = (int)tot;
    st->n_scan = (int)tot;
  }
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < 8; r++) {
    if ((bal[r] >> lane) & 1u) {
      const uint32_t o = pos + __popc(bal[r] & lt);
      const float4 q = make_float4(pt[r].x, pt[r].y, pt[r].z, t_rel);
      cloud[o] = q;
      orig[o] = q;
      if (in_map)
        in_map[o] = tf_apply(T, q);
    }
    pos += __popc(bal[r]);
  }
}
__global__ void k_scanstate_zero(ScanState *__restrict__ st) {
  st->n_raw = st->n_orig = st->n_scan = 0;
  st->vox_n = st->vox_passthrough = 0;
}

__global__ void __launch_bounds__(256) k_transform(const float4 *__restrict__ in, const int *__restrict__ d_n, Tf32 T,
                                                  float4 *__restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_n)
    return;
  out[i] = tf_apply(T, in[i]);
}

// ---- LidarHelper::remove_motion_blur loop body, LidarHelper.cpp:71-85 ---------------------------------
// transformPoint(p, Affine3f): ((r0*x + r1*y) + r2*z) + t. Knots: n_knots 3x4 transforms; point i sits at sweep
// fraction 1 - (total - i)/total = i/total of [t - total * raw_point_dt, t]; with n_knots == total every point has its own
// transform (the reference's case), otherwise the two bracketing knots are blended linearly (documented approximation).
__global__ void __launch_bounds__(256) k_deskew(float4 *__restrict__ cloud, const int *__restrict__ d_n, const float *__restrict__ knots,
                                               int n_knots, int total) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_n)
    return;
  float T[12];
  if (n_knots == total) {
#pragma unroll
    for (int k = 0; k < 12; k++)
      T[k] = knots[(size_t)i * 12 + k];
  } else {
    float u = (float)i / (float)total * (float)(n_knots - 1); // point i sits (total - i) * raw_point_dt before the scan time (:68)
    int k0 = min((int)u, n_knots - 2);
    float a = u - (float)k0;
#pragma unroll
    for (int k = 0; k < 12; k++)
      T[k] = (1.f - a) * knots[(size_t)k0 * 12 + k] + a * knots[(size_t)(k0 + 1) * 12 + k];
  }
  float4 p = cloud[i];
  float4 q;
  q.x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[0], p.x), __fmul_rn(T[1], p.y)), __fmul_rn(T[2], p.z)), T[3]);
  q.y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[4], p.x), __fmul_rn(T[5], p.y)), __fmul_rn(T[6], p.z)), T[7]);
  q.z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[8], p.x), __fmul_rn(T[9], p.y)), __fmul_rn(T[10], p.z)), T[11]);
  q.w = p.w;
  cloud[i] = q;
}

// ---- pcl::VoxelGrid<PointXYZI>, LidarHelper.cpp:89-101,214 -------------------------------------------
// PCL 1.10 filters/impl/voxel_grid.hpp: bounds -> min_b/div_b -> linear voxel index -> sort -> centroids in
// ascending index. Here: bbox reduce, voxel index, stable radix sort by index (ties in input order), run heads, one thread
// per voxel sums its points sequentially in fp32 and divides by the count (kernel list below).
struct VoxGridDesc {
  int minb[3];
  int divb[3];
  int passthrough; // PCL: "Leaf size is too small" -> output = input
  float inv_leaf;
};
__device__ __forceinline__ void voxgrid_setup_dev(const uint32_t *bbox, int n, float leaf, VoxGridDesc *__restrict__ vg) {
  const float inv = __fdiv_rn(1.0f, leaf);
  vg->inv_leaf = inv;
  if (n <= 0) {
    vg->passthrough = 0;
    for (int a = 0; a < 3; a++)
      vg->minb[a] = 0, vg->divb[a] = 1;
    return;
  }
  long long d[3];
  for (int a = 0; a < 3; a++) {
    const float mn = ord2f(bbox[a]), mx = ord2f(bbox[3 + a]);
    d[a] = (long long)(__fmul_rn(__fsub_rn(mx, mn), inv)) + 1;
    vg->minb[a] = (int)floorf(__fmul_rn(mn, inv));
    int maxb = (int)floorf(__fmul_rn(mx, inv));
    vg->divb[a] = maxb - vg->minb[a] + 1;
  }
  vg->passthrough = (d[0] * d[1] * d[2] > 2147483647ll) ? 1 : 0;
}

// The voxel filter in 7 launches (was 15 with a 2-kernel scan, 22 with the 3-kernel radix passes):
//   k_bbox_voxgrid     min/max of the cloud; the CTA that finishes last derives the voxel grid (PCL's getMinMax3D + leaf
//                      arithmetic) and puts `bbox` back into its initial state (invariant: initial between uses)
//   k_vox_index_hist   voxel index per point (sort key) + the digit histograms of all four radix passes
//   4 x k_radix_pass   stable sort by voxel index (prims.cuh)
//   k_vox_reduce       (update.cuh) run heads of the sorted keys, their stable single-pass ranking (look-back), one centroid per run
//                      summed in input order, the new count; or the unchanged copy when PCL would refuse the leaf size
__global__ void __launch_bounds__(256) k_bbox_voxgrid(const float4 *__restrict__ pts, const int *__restrict__ d_n, uint32_t *bbox,
                                                     unsigned int *done, float leaf, VoxGridDesc *__restrict__ vg) {
  __shared__ float smn[8][3], smx[8][3];
  __shared__ uint32_t s_last;
  const int n = *d_n;
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 p = pts[i];
    mn[0] = fminf(mn[0], p.x), mx[0] = fmaxf(mx[0], p.x);
    mn[1] = fminf(mn[1], p.y), mx[1] = fmaxf(mx[1], p.y);
    mn[2] = fminf(mn[2], p.z), mx[2] = fmaxf(mx[2], p.z);
  }
#pragma unroll
  for (int a = 0; a < 3; a++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) {
#pragma unroll
    for (int a = 0; a < 3; a++)
      smn[wid][a] = mn[a], smx[wid][a] = mx[a];
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int a = threadIdx.x;
    float lo = INFINITY, hi = -INFINITY;
    for (int w = 0; w < 8; w++)
      lo = fminf(lo, smn[w][a]), hi = fmaxf(hi, smx[w][a]);
    if (lo <= hi) {
      atomicMin(&bbox[a], f2ord(lo));
      atomicMax(&bbox[3 + a], f2ord(hi));
    }
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0)
    s_last = atomicAdd(done, 1u) == gridDim.x - 1u;
  __syncthreads();
  if (!s_last || threadIdx.x != 0)
    return;
  __threadfence();
  uint32_t bb[6];
  for (int c = 0; c < 6; c++) {
    bb[c] = *(volatile uint32_t *)&bbox[c];
    bbox[c] = c < 3 ? 0xffffffffu : 0u;
  }
  voxgrid_setup_dev(bb, n, leaf, vg);
  *done = 0u;
}

__global__ void __launch_bounds__(RS_BLOCK) k_vox_index_hist(const float4 *__restrict__ cloud, const int *__restrict__ d_n,
                                                            const VoxGridDesc *__restrict__ vgd, uint32_t *__restrict__ keys,
                                                            uint32_t *__restrict__ vals, uint32_t *ghist, uint32_t *__restrict__ dbase,
                                                            unsigned int *done) {
  __shared__ uint32_t h[RS_PASSES][256];
  const int n = *d_n;
  const uint32_t active = n > 0 ? (uint32_t)((n + RS_TILE - 1) / RS_TILE) : 1u;
  if (blockIdx.x >= active)
    return;
#pragma unroll
  for (int p = 0; p < RS_PASSES; p++)
    h[p][threadIdx.x] = 0;
  __syncthreads();
  const VoxGridDesc vg = *vgd;
  const int base = blockIdx.x * RS_TILE;
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const int i = base + k * RS_BLOCK + threadIdx.x;
    if (i < n) {
      const float4 p = cloud[i];
      const int i0 = (int)floorf(__fmul_rn(p.x, vg.inv_leaf)) - vg.minb[0];
      const int i1 = (int)floorf(__fmul_rn(p.y, vg.inv_leaf)) - vg.minb[1];
      const int i2 = (int)floorf(__fmul_rn(p.z, vg.inv_leaf)) - vg.minb[2];
      const uint32_t key = (uint32_t)(i0 + i1 * vg.divb[0] + i2 * vg.divb[0] * vg.divb[1]);
      keys[i] = key;
      vals[i] = (uint32_t)i;
      radix_hist4_add(h, key);
    }
  }
  radix_hist4_publish(h, ghist, dbase, done, active);
}

} // namespace ml
