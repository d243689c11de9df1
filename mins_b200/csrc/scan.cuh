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
// ascending index. Here: bbox reduce, k_vox_index, stable radix sort by index (ties in input order), segment
// heads, one thread per voxel sums its points sequentially in fp32 and divides by the count.
struct VoxGridDesc {
  int minb[3];
  int divb[3];
  int passthrough; // PCL: "Leaf size is too small" -> output = input
  float inv_leaf;
};
__global__ void k_voxgrid_setup(const uint32_t *__restrict__ bbox, const int *__restrict__ d_n, float leaf, VoxGridDesc *__restrict__ vg) {
  if (threadIdx.x != 0 || blockIdx.x != 0)
    return;
  const float inv = __fdiv_rn(1.0f, leaf);
  vg->inv_leaf = inv;
  if (*d_n <= 0) {
    vg->passthrough = 0;
    for (int a = 0; a < 3; a++)
      vg->minb[a] = 0, vg->divb[a] = 1;
    return;
  }
  long long d[3];
  for (int a = 0; a < 3; a++) {
    float mn = __uint_as_float(0), mx = 0;
    uint32_t u0 = bbox[a], u1 = bbox[3 + a];
    mn = __uint_as_float((u0 & 0x80000000u) ? (u0 & 0x7fffffffu) : ~u0);
    mx = __uint_as_float((u1 & 0x80000000u) ? (u1 & 0x7fffffffu) : ~u1);
    d[a] = (long long)(__fmul_rn(__fsub_rn(mx, mn), inv)) + 1;
    vg->minb[a] = (int)floorf(__fmul_rn(mn, inv));
    int maxb = (int)floorf(__fmul_rn(mx, inv));
    vg->divb[a] = maxb - vg->minb[a] + 1;
  }
  vg->passthrough = (d[0] * d[1] * d[2] > 2147483647ll) ? 1 : 0;
}
__global__ void __launch_bounds__(256) k_vox_index(const float4 *__restrict__ cloud, const int *__restrict__ d_n,
                                                  const VoxGridDesc *__restrict__ vgd, uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_n)
    return;
  const VoxGridDesc vg = *vgd;
  float4 p = cloud[i];
  int i0 = (int)floorf(__fmul_rn(p.x, vg.inv_leaf)) - vg.minb[0];
  int i1 = (int)floorf(__fmul_rn(p.y, vg.inv_leaf)) - vg.minb[1];
  int i2 = (int)floorf(__fmul_rn(p.z, vg.inv_leaf)) - vg.minb[2];
  keys[i] = (uint32_t)(i0 + i1 * vg.divb[0] + i2 * vg.divb[0] * vg.divb[1]);
  vals[i] = (uint32_t)i;
}
// head flag of each run of equal keys in the sorted order
__global__ void __launch_bounds__(256) k_seg_heads(const uint32_t *__restrict__ keys, const int *__restrict__ d_n, uint32_t *__restrict__ head) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_n)
    return;
  head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}
// one thread per sorted element that is a head: sum the run sequentially (fp32, input order inside the voxel)
__global__ void __launch_bounds__(256) k_vox_centroid(const float4 *__restrict__ cloud, const uint32_t *__restrict__ keys,
                                                     const uint32_t *__restrict__ vals, const uint32_t *__restrict__ head,
                                                     const uint32_t *__restrict__ pos, const int *__restrict__ d_n,
                                                     float4 *__restrict__ out) {
  const int n = *d_n;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !head[i])
    return;
  const uint32_t key = keys[i];
  float sx = 0.f, sy = 0.f, sz = 0.f, sw = 0.f;
  int j = i;
  for (; j < n && keys[j] == key; j++) {
    float4 p = cloud[vals[j]];
    sx = __fadd_rn(sx, p.x), sy = __fadd_rn(sy, p.y), sz = __fadd_rn(sz, p.z), sw = __fadd_rn(sw, p.w);
  }
  float c = (float)(j - i);
  out[pos[i]] = make_float4(__fdiv_rn(sx, c), __fdiv_rn(sy, c), __fdiv_rn(sz, c), __fdiv_rn(sw, c));
}
// publish the down-sampled count (or copy through when PCL would refuse the leaf size)
__global__ void k_vox_finish(const VoxGridDesc *__restrict__ vg, const uint32_t *__restrict__ pos, int *__restrict__ d_n_inout,
                             int *__restrict__ d_passthrough) {
  if (threadIdx.x != 0 || blockIdx.x != 0)
    return;
  *d_passthrough = vg->passthrough;
  if (!vg->passthrough)
    *d_n_inout = (int)pos[*d_n_inout];
}
__global__ void __launch_bounds__(256) k_copy_if_flag(const float4 *__restrict__ src, float4 *__restrict__ dst, const int *__restrict__ d_n,
                                                     const int *__restrict__ d_flag, int want) {
  if ((*d_flag != 0) != (want != 0))
    return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < *d_n)
    dst[i] = src[i];
}

} // namespace ml
