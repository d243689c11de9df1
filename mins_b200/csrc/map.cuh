// Local-map kernels: dense-grid build (replaces KD_TREE::Build), voxel-deduplicating insert (replaces
// KD_TREE::Add_Points), re-anchoring (replaces LidarHelper::propagate_map_to_newest_clone's point work).
// Reference citations relative to /root/reference/mins/src/update/lidar/.
#pragma once
#include "prims.cuh"

namespace ml {

// order-preserving float <-> uint mapping for atomicMin/Max on floats
__device__ __forceinline__ uint32_t f2ord(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// bbox[0..2] = min (ordered uint), bbox[3..5] = max
__global__ void k_bbox_init(uint32_t *bbox) {
  if (threadIdx.x < 3)
    bbox[threadIdx.x] = 0xffffffffu;
  else if (threadIdx.x < 6)
    bbox[threadIdx.x] = 0u;
}

__global__ void __launch_bounds__(256) k_bbox(const float4 *__restrict__ pts, const int *__restrict__ d_n, uint32_t *__restrict__ bbox) {
  const int n = *d_n;
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float4 p = pts[i];
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
  // one set of 6 atomics per CTA (per-warp atomics on 6 addresses serialise: 19 us for a 131 072-point scan)
  __shared__ float smn[8][3], smx[8][3];
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
    for (int w = 0; w < (int)(blockDim.x >> 5); w++)
      lo = fminf(lo, smn[w][a]), hi = fmaxf(hi, smx[w][a]);
    if (lo <= hi) {
      atomicMin(&bbox[a], f2ord(lo));
      atomicMax(&bbox[3 + a], f2ord(hi));
    }
  }
}

// one thread: choose origin/dims; the cell edge doubles until the grid fits max_cells.
__global__ void k_grid_setup(const uint32_t *__restrict__ bbox, const int *__restrict__ d_n, double cell0, int max_cells,
                             GridDesc *__restrict__ g) {
  if (threadIdx.x != 0 || blockIdx.x != 0)
    return;
  const int n = *d_n;
  g->n_pts = n;
  if (n <= 0) {
    g->valid = 0;
    g->ncells = 0;
    g->dimx = g->dimy = g->dimz = 0;
    g->cell = cell0;
    g->inv_cell = 1.0 / cell0;
    g->ox = g->oy = g->oz = 0;
    return;
  }
  double mn[3], mx[3];
  for (int a = 0; a < 3; a++) {
    mn[a] = (double)ord2f(bbox[a]);
    mx[a] = (double)ord2f(bbox[3 + a]);
  }
  double cell = cell0;
  for (;;) {
    double o[3];
    long long d[3];
    for (int a = 0; a < 3; a++) {
      o[a] = floor(mn[a] / cell) * cell - cell; // one empty cell of padding
      d[a] = (long long)floor((mx[a] - o[a]) / cell) + 2;
    }
    if (d[0] * d[1] * d[2] <= (long long)max_cells) {
      g->ox = o[0], g->oy = o[1], g->oz = o[2];
      g->dimx = (int)d[0], g->dimy = (int)d[1], g->dimz = (int)d[2];
      g->ncells = (int)(d[0] * d[1] * d[2]);
      break;
    }
    cell *= 2.0;
  }
  g->cell = cell;
  g->inv_cell = 1.0 / cell;
  g->valid = 1;
}

__global__ void __launch_bounds__(256) k_zero_cells(uint32_t *__restrict__ cnt, const GridDesc *__restrict__ g) {
  const int n = g->ncells + 1;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    cnt[i] = 0;
}

__device__ __forceinline__ int cell_of(const GridDesc &g, float x, float y, float z) {
  int cx = cell_coord((double)x, g.ox, g.inv_cell);
  int cy = cell_coord((double)y, g.oy, g.inv_cell);
  int cz = cell_coord((double)z, g.oz, g.inv_cell);
  cx = min(max(cx, 0), g.dimx - 1);
  cy = min(max(cy, 0), g.dimy - 1);
  cz = min(max(cz, 0), g.dimz - 1);
  return (cz * g.dimy + cy) * g.dimx + cx;
}

// cell id and arrival rank of each map point (rank order inside a cell is irrelevant: the kNN comparator is a
// total order on (d2, map index))
__global__ void __launch_bounds__(256) k_cell_count(const float4 *__restrict__ pts, const GridDesc *__restrict__ gd,
                                                   uint32_t *__restrict__ cnt, uint32_t *__restrict__ pt_cell,
                                                   uint32_t *__restrict__ pt_rank) {
  const GridDesc g = *gd;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.n_pts)
    return;
  float4 p = pts[i];
  int c = cell_of(g, p.x, p.y, p.z);
  pt_cell[i] = (uint32_t)c;
  pt_rank[i] = atomicAdd(&cnt[c], 1u);
}

__global__ void __launch_bounds__(256) k_cell_scatter(const float4 *__restrict__ pts, const GridDesc *__restrict__ gd,
                                                     const uint32_t *__restrict__ cell_start, const uint32_t *__restrict__ pt_cell,
                                                     const uint32_t *__restrict__ pt_rank, float4 *__restrict__ cell_pts) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= gd->n_pts)
    return;
  float4 p = pts[i];
  p.w = __int_as_float(i);
  cell_pts[cell_start[pt_cell[i]] + pt_rank[i]] = p;
}

// ------------------------------------------------------------------------------------------------
// Insert with per-voxel de-duplication — KD_TREE::Add_Points(points, true), ikd_Tree.cpp:401-479.
// Sequential reference semantics folded into an order-independent rule (DESIGN.md §map insert):
// per touched voxel the survivor is argmin over {existing points inside the voxel box, new points of the voxel}
// of (fp32 squared distance to the box centre, later-in-sequence first); every other candidate of the box dies.
// Sequence number = map index for existing points, M + i for the i-th new point.
// ------------------------------------------------------------------------------------------------
struct VoxSlot {
  unsigned long long key;  // 0 = empty; else packed voxel coordinates + 1
  unsigned long long best; // (dist bits << 32) | (0xffffffff - seq), atomicMin
};

__device__ __forceinline__ unsigned long long vox_key(float x, float y, float z, float ds, int *kx, int *ky, int *kz) {
  // floor(x / ds) exactly as ikd_Tree.cpp:411 (fp32 division)
  float fx = floorf(__fdiv_rn(x, ds)), fy = floorf(__fdiv_rn(y, ds)), fz = floorf(__fdiv_rn(z, ds));
  *kx = (int)fx, *ky = (int)fy, *kz = (int)fz;
  const unsigned long long B = 1ull << 20;
  return ((((unsigned long long)(*kx + (long long)B)) & 0x1fffffull) << 42) | ((((unsigned long long)(*ky + (long long)B)) & 0x1fffffull) << 21) |
         (((unsigned long long)(*kz + (long long)B)) & 0x1fffffull) | (1ull << 63);
}
__device__ __forceinline__ uint32_t hash64(unsigned long long k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return (uint32_t)k;
}
// box centre and membership test of ikd_Tree.cpp:411-419,1050-1053 from the integer voxel coordinate
__device__ __forceinline__ void vox_box(int k, float ds, float *bmin, float *bmax, float *mid) {
  *bmin = __fmul_rn((float)k, ds);
  *bmax = __fadd_rn(*bmin, ds);
  *mid = (float)((double)*bmin + ((double)__fsub_rn(*bmax, *bmin)) / 2.0);
}

__global__ void __launch_bounds__(256) k_vox_clear(VoxSlot *__restrict__ tab, int nslots) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nslots; i += gridDim.x * blockDim.x) {
    tab[i].key = 0ull;
    tab[i].best = ~0ull;
  }
}

// new points claim their voxel slot and compete
__global__ void __launch_bounds__(256) k_vox_insert_new(const float4 *__restrict__ add, const int *__restrict__ d_nadd,
                                                       const int *__restrict__ d_M, float ds, VoxSlot *__restrict__ tab,
                                                       uint32_t mask, uint32_t *__restrict__ add_slot) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_nadd)
    return;
  const int M = *d_M;
  float4 p = add[i];
  int kx, ky, kz;
  unsigned long long key = vox_key(p.x, p.y, p.z, ds, &kx, &ky, &kz);
  float bmin, bmax, mx, my, mz;
  vox_box(kx, ds, &bmin, &bmax, &mx);
  vox_box(ky, ds, &bmin, &bmax, &my);
  vox_box(kz, ds, &bmin, &bmax, &mz);
  float d = dist2_ref(p.x, p.y, p.z, mx, my, mz);
  uint32_t s = hash64(key) & mask;
  for (;;) {
    unsigned long long prev = atomicCAS(&tab[s].key, 0ull, key);
    if (prev == 0ull || prev == key)
      break;
    s = (s + 1) & mask;
  }
  add_slot[i] = s;
  unsigned long long cand = ((unsigned long long)__float_as_uint(d) << 32) | (unsigned long long)(0xffffffffu - (uint32_t)(M + i));
  atomicMin(&tab[s].best, cand);
}

// existing points inside a touched voxel box compete; contender[m] = slot+1 or 0
__global__ void __launch_bounds__(256) k_vox_compete_old(const float4 *__restrict__ map, const int *__restrict__ d_M, float ds,
                                                        VoxSlot *__restrict__ tab, uint32_t mask, uint32_t *__restrict__ contender) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= *d_M)
    return;
  float4 p = map[m];
  int kx, ky, kz;
  unsigned long long key = vox_key(p.x, p.y, p.z, ds, &kx, &ky, &kz);
  uint32_t s = hash64(key) & mask;
  uint32_t found = 0;
  for (;;) {
    unsigned long long k = tab[s].key;
    if (k == key) {
      found = s + 1;
      break;
    }
    if (k == 0ull)
      break;
    s = (s + 1) & mask;
  }
  if (found) {
    float bx0, bx1, by0, by1, bz0, bz1, mx, my, mz;
    vox_box(kx, ds, &bx0, &bx1, &mx);
    vox_box(ky, ds, &by0, &by1, &my);
    vox_box(kz, ds, &bz0, &bz1, &mz);
    bool in = bx0 <= p.x && bx1 > p.x && by0 <= p.y && by1 > p.y && bz0 <= p.z && bz1 > p.z;
    if (in) {
      float d = dist2_ref(p.x, p.y, p.z, mx, my, mz);
      unsigned long long cand = ((unsigned long long)__float_as_uint(d) << 32) | (unsigned long long)(0xffffffffu - (uint32_t)m);
      atomicMin(&tab[found - 1].best, cand);
    } else {
      found = 0;
    }
  }
  contender[m] = found;
}

// keep[] over the concatenation [existing (M) | new (nadd)]; d_total = M + nadd
__global__ void __launch_bounds__(256) k_vox_flags(const int *__restrict__ d_M, const int *__restrict__ d_nadd,
                                                  const VoxSlot *__restrict__ tab, const uint32_t *__restrict__ contender,
                                                  const uint32_t *__restrict__ add_slot, int downsample, uint32_t *__restrict__ keep,
                                                  int *__restrict__ d_total) {
  const int M = *d_M, na = *d_nadd;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0)
    *d_total = M + na;
  if (i >= M + na)
    return;
  uint32_t k = 1;
  if (downsample) {
    if (i < M) {
      uint32_t c = contender[i];
      if (c) {
        uint32_t seq = 0xffffffffu - (uint32_t)(tab[c - 1].best & 0xffffffffull);
        k = (seq == (uint32_t)i) ? 1u : 0u;
      }
    } else {
      uint32_t seq = 0xffffffffu - (uint32_t)(tab[add_slot[i - M]].best & 0xffffffffull);
      k = (seq == (uint32_t)i) ? 1u : 0u;
    }
  }
  keep[i] = k;
}

// stable compaction of [map | add] into out using the exclusive scan of keep
__global__ void __launch_bounds__(256) k_compact2(const float4 *__restrict__ a, const int *__restrict__ d_na,
                                                 const float4 *__restrict__ b, const int *__restrict__ d_nb,
                                                 const uint32_t *__restrict__ keep, const uint32_t *__restrict__ pos,
                                                 float4 *__restrict__ out, int cap, int *__restrict__ d_overflow) {
  const int na = *d_na, nb = d_nb ? *d_nb : 0;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= na + nb)
    return;
  if (keep[i]) {
    uint32_t p = pos[i];
    if ((int)p < cap)
      out[p] = i < na ? a[i] : b[i - na];
    else
      atomicExch(d_overflow, 1);
  }
}

// ------------------------------------------------------------------------------------------------
// Re-anchor — LidarHelper.cpp:219-226 (decay filters) and :261 (pcl::transformPointCloud with a Matrix4d:
// float(((m0*x + m1*y) + m2*z) + m3) in fp64). keep flags + transformed copy.
// ------------------------------------------------------------------------------------------------
struct ReanchorArgs {
  double T[16];
  double now_minus_FT, decay_time, decay_dist;
};
__global__ void __launch_bounds__(256) k_reanchor(const float4 *__restrict__ in, const int *__restrict__ d_n, ReanchorArgs a,
                                                 float4 *__restrict__ out, uint32_t *__restrict__ keep) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_n)
    return;
  float4 p = in[i];
  bool k = !((double)p.w + a.decay_time < a.now_minus_FT);
  k = k && !((double)norm3_ref(p.x, p.y, p.z) > a.decay_dist);
  double x = p.x, y = p.y, z = p.z;
  float4 q;
  q.x = (float)__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(a.T[0], x), __dmul_rn(a.T[1], y)), __dmul_rn(a.T[2], z)), a.T[3]);
  q.y = (float)__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(a.T[4], x), __dmul_rn(a.T[5], y)), __dmul_rn(a.T[6], z)), a.T[7]);
  q.z = (float)__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(a.T[8], x), __dmul_rn(a.T[9], y)), __dmul_rn(a.T[10], z)), a.T[11]);
  q.w = p.w;
  out[i] = q;
  keep[i] = k ? 1u : 0u;
}

__global__ void k_set_int(int *p, int v) { *p = v; }
__global__ void k_copy_int(int *dst, const int *src) { *dst = *src; }

} // namespace ml
