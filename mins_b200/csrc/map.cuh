// Local-map kernels: dense-grid build (replaces KD_TREE::Build), voxel-deduplicating insert (replaces
// KD_TREE::Add_Points), re-anchoring (replaces LidarHelper::propagate_map_to_newest_clone's point work).
// Reference citations relative to /root/reference/mins/src/update/lidar/.
#pragma once
#include "prims.cuh"

namespace ml {

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
__device__ __forceinline__ void grid_setup_dev(const uint32_t *bbox, int n, double cell0, int max_cells, GridDesc *g) {
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
__global__ void k_grid_setup(const uint32_t *__restrict__ bbox, const int *__restrict__ d_n, double cell0, int max_cells,
                             GridDesc *__restrict__ g) {
  if (threadIdx.x == 0 && blockIdx.x == 0)
    grid_setup_dev(bbox, *d_n, cell0, max_cells, g);
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

// new points claim their voxel slot and compete. XF: the new points are pointcloud_original still in the sensor frame;
// transform them (LidarHelper.cpp:203) on the way and leave the transformed copy in `add` for the merge.
template <bool XF>
__global__ void __launch_bounds__(256) k_vox_insert_new(const float4 *__restrict__ src, Tf32 T, float4 *__restrict__ add,
                                                       const int *__restrict__ d_nadd, const int *__restrict__ d_M, float ds,
                                                       VoxSlot *__restrict__ tab, uint32_t mask, uint32_t *__restrict__ add_slot) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_nadd)
    return;
  const int M = *d_M;
  float4 p;
  if (XF) {
    p = tf_apply(T, src[i]);
    add[i] = p;
  } else {
    p = add[i];
  }
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

// ------------------------------------------------------------------------------------------------
// One-pass map merge: survivors of [a (na) | b (nb)] stably compacted into `out`, the bounding box of the survivors, and -
// by the CTA that finishes last - the new point count, the kNN grid descriptor and a result block in pinned host memory.
// Replaces keep-flag, 2-kernel scan, compaction, count copy, bbox init, bbox and grid set-up launches (8 -> 1) and two of the
// three passes over the map. Survivor rule by MODE:
//   MG_ALL   every point (Add_Points without down-sampling)
//   MG_VOX   the per-voxel contest of ikd_Tree.cpp:401-479 decided in `tab` (see above)
//   MG_FLAGS keep[i] != 0 (re-anchor: decay filters)
// Stable order = decoupled look-back over MG_TILE-point tiles (prims.cuh). `bbox` must hold the k_bbox_init values on entry.
// ------------------------------------------------------------------------------------------------
enum { MG_ALL = 0, MG_VOX = 1, MG_FLAGS = 2 };
constexpr int MG_BLOCK = 256, MG_ROWS = 8, MG_TILE = MG_BLOCK * MG_ROWS; // 2048-point tiles (4 / 8 / 16 rows measured: 31 -> 21 / 21 / 24 us)
struct MergeResult { // pinned host memory, written by the last CTA
  int n_in, n_out, overflow, ncells;
};
template <int MODE>
__global__ void __launch_bounds__(MG_BLOCK)
    k_map_merge(const float4 *__restrict__ a, const int *d_na, const float4 *__restrict__ b, const int *__restrict__ d_nb,
                const VoxSlot *__restrict__ tab, const uint32_t *__restrict__ contender, const uint32_t *__restrict__ add_slot,
                const uint32_t *__restrict__ keep, float4 *__restrict__ out, int cap, unsigned long long *tile_state,
                unsigned int *tile_ctr, uint32_t epoch, uint32_t *bbox, unsigned int *done_ctr,
                int *d_total, double cell0, int max_cells, GridDesc *g, int *d_map_n, MergeResult *res) {
  __shared__ uint32_t wsum[MG_BLOCK / 32];
  __shared__ uint32_t s_tile, s_prefix, s_last;
  __shared__ float smn[MG_BLOCK / 32][3], smx[MG_BLOCK / 32][3];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int na = *d_na, nb = d_nb ? *d_nb : 0, n = na + nb;
  const uint32_t active = n > 0 ? (uint32_t)((n + MG_TILE - 1) / MG_TILE) : 1u;
  if (blockIdx.x < active) { // CTA-uniform; the launch is sized by a host-side upper bound of n
    if (threadIdx.x == 0)
      s_tile = lb_ticket(tile_ctr, active);
    __syncthreads();
    const uint32_t tile = s_tile;
    const long long tbase = (long long)tile * MG_TILE;
    const int wbase = (int)tbase + wid * (MG_TILE / (MG_BLOCK / 32));
    // three unrolled phases so that the loads of all rows are in flight together (a ballot per row in the same loop would
    // serialise them: 16 dependent round trips instead of 2)
    float4 pt[MG_ROWS];
    uint32_t sel[MG_ROWS]; // MG_VOX: voxel slot + 1 whose winner decides (0: not contested); MG_FLAGS: keep flag
    uint32_t bal[MG_ROWS];
    uint32_t cnt = 0;
#pragma unroll
    for (int r = 0; r < MG_ROWS; r++) {
      const int i = wbase + r * 32 + lane;
      sel[r] = 0u;
      if (i < n) {
        pt[r] = i < na ? a[i] : b[i - na];
        if (MODE == MG_VOX)
          sel[r] = i < na ? contender[i] : add_slot[i - na] + 1u;
        else if (MODE == MG_FLAGS)
          sel[r] = keep[i];
      }
    }
    if (MODE == MG_VOX) {
#pragma unroll
      for (int r = 0; r < MG_ROWS; r++) {
        const int i = wbase + r * 32 + lane;
        // the voxel's winner is the sequence number in the low word of `best`; uncontested points win by default
        const uint32_t win = sel[r] ? 0xffffffffu - (uint32_t)(tab[sel[r] - 1u].best & 0xffffffffull) : (uint32_t)i;
        sel[r] = win == (uint32_t)i ? 1u : 0u;
      }
    }
#pragma unroll
    for (int r = 0; r < MG_ROWS; r++) {
      const int i = wbase + r * 32 + lane;
      const bool k = i < n && (MODE == MG_ALL || sel[r] != 0u);
      bal[r] = __ballot_sync(0xffffffffu, k);
      cnt += __popc(bal[r]);
    }
    if (lane == 0)
      wsum[wid] = cnt;
    __syncthreads();
    if (wid == 0) {
      uint32_t tot = lane < MG_BLOCK / 32 ? wsum[lane] : 0u;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
        tot += __shfl_xor_sync(0xffffffffu, tot, o);
      const uint32_t excl = lb_exclusive_prefix(tile_state, tile, epoch, tot);
      if (lane == 0) {
        s_prefix = excl;
        if (tbase + MG_TILE >= (long long)n) // the tile that holds the end of the input knows the total
          *d_total = (int)(excl + tot);
      }
    }
    __syncthreads();
    uint32_t pos = s_prefix;
    for (int w = 0; w < wid; w++)
      pos += wsum[w];
    const uint32_t lt = (1u << lane) - 1u;
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int r = 0; r < MG_ROWS; r++) {
      if ((bal[r] >> lane) & 1u) {
        const uint32_t o = pos + __popc(bal[r] & lt);
        if ((int)o < cap)
          out[o] = pt[r];
        mn[0] = fminf(mn[0], pt[r].x), mx[0] = fmaxf(mx[0], pt[r].x);
        mn[1] = fminf(mn[1], pt[r].y), mx[1] = fmaxf(mx[1], pt[r].y);
        mn[2] = fminf(mn[2], pt[r].z), mx[2] = fmaxf(mx[2], pt[r].z);
      }
      pos += __popc(bal[r]);
    }
#pragma unroll
    for (int c = 0; c < 3; c++) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
        mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
      }
    }
    if (lane == 0) {
#pragma unroll
      for (int c = 0; c < 3; c++)
        smn[wid][c] = mn[c], smx[wid][c] = mx[c];
    }
    __syncthreads();
    if (threadIdx.x < 3) {
      const int c = threadIdx.x;
      float lo = INFINITY, hi = -INFINITY;
      for (int w = 0; w < MG_BLOCK / 32; w++)
        lo = fminf(lo, smn[w][c]), hi = fmaxf(hi, smx[w][c]);
      if (lo <= hi) {
        atomicMin(&bbox[c], f2ord(lo));
        atomicMax(&bbox[3 + c], f2ord(hi));
      }
    }
  }
  // last CTA out: new count, grid descriptor, result block
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0)
    s_last = atomicAdd(done_ctr, 1u) == gridDim.x - 1u;
  __syncthreads();
  if (!s_last || threadIdx.x != 0)
    return;
  __threadfence();
  *done_ctr = 0u;
  const int total = *(volatile int *)d_total;
  const bool ovf = total > cap;
  if (!ovf) {
    *d_map_n = total;
    uint32_t bb[6];
    for (int c = 0; c < 6; c++)
      bb[c] = *(volatile uint32_t *)&bbox[c];
    grid_setup_dev(bb, total, cell0, max_cells, g);
  } else { // the map is double-buffered: the old one is intact; the grid is rebuilt for it by the host's error path
    grid_setup_dev(bbox, 0, cell0, max_cells, g);
  }
  res->n_in = n;
  res->n_out = total;
  res->overflow = ovf ? 1 : 0;
  res->ncells = g->ncells;
  __threadfence_system();
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
