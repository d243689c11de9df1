// The per-scan measurement-update kernels: query binning, exact grid kNN, per-point plane/residual/Jacobian/
// chi2 rows, fixed-order reduction to the information form. Reference: mins/src/update/lidar/UpdaterLidar.cpp:266-397,
// LidarHelper.cpp:268-328, mins/src/state/StateHelper.cpp:645-680, mins/src/update/UpdaterStatistics.cpp:94-117.
#pragma once
#include "prims.cuh"
#include "map.cuh"
#include "scan.cuh"

namespace ml {

// ------------------------------------------------------------------------------------------------
// Query binning: queries are grouped by grid cell with a hash table (O(N), no pass over the grid).
// ------------------------------------------------------------------------------------------------
struct QSlot {
  unsigned long long key; // 0 = empty
  uint32_t cnt;
  uint32_t qbase;
};
__device__ __forceinline__ unsigned long long cell_key(int cx, int cy, int cz) {
  const long long B = 1ll << 20;
  return ((((unsigned long long)(cx + B)) & 0x1fffffull) << 42) | ((((unsigned long long)(cy + B)) & 0x1fffffull) << 21) |
         (((unsigned long long)(cz + B)) & 0x1fffffull) | (1ull << 63);
}
__device__ __forceinline__ void cell_unkey(unsigned long long k, int *cx, int *cy, int *cz) {
  const int B = 1 << 20;
  *cx = (int)((k >> 42) & 0x1fffffull) - B;
  *cy = (int)((k >> 21) & 0x1fffffull) - B;
  *cz = (int)(k & 0x1fffffull) - B;
}

// Places one query into the hash table. Called by all lanes of `act` together. Warp-aggregated: consecutive rays mostly
// share a cell, so one lane per distinct key probes the table and reserves the ranks of all its peers with one atomicAdd.
__device__ __forceinline__ void bin_query(const float4 q, const unsigned act, const GridDesc *__restrict__ gd, QSlot *__restrict__ tab,
                                          const uint32_t mask, uint32_t *slot_out, uint32_t *rank_out) {
  const double ox = gd->ox, oy = gd->oy, oz = gd->oz, inv = gd->inv_cell;
  // clamp far-away queries to a band around the grid so the key packing cannot overflow; such queries have no
  // candidate within reach anyway
  const int lim = 1 << 19;
  const int cx = max(-lim, min(lim, cell_coord((double)q.x, ox, inv)));
  const int cy = max(-lim, min(lim, cell_coord((double)q.y, oy, inv)));
  const int cz = max(-lim, min(lim, cell_coord((double)q.z, oz, inv)));
  const unsigned long long key = cell_key(cx, cy, cz);
  const int lane = threadIdx.x & 31;
  const unsigned peers = __match_any_sync(act, key);
  const int leader = __ffs(peers) - 1;
  uint32_t s = 0, base = 0;
  if (lane == leader) {
    s = hash64(key) & mask;
    for (;;) {
      unsigned long long prev = atomicCAS(&tab[s].key, 0ull, key);
      if (prev == 0ull || prev == key)
        break;
      s = (s + 1) & mask;
    }
    base = atomicAdd(&tab[s].cnt, (uint32_t)__popc(peers));
  }
  s = __shfl_sync(peers, s, leader);
  base = __shfl_sync(peers, base, leader);
  *slot_out = s;
  *rank_out = base + (uint32_t)__popc(peers & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(256) k_query_bin(const float4 *__restrict__ q_in_map, const ScanState *__restrict__ st,
                                                  const GridDesc *__restrict__ gd, QSlot *__restrict__ tab, uint32_t mask,
                                                  uint32_t *__restrict__ q_slot, uint32_t *__restrict__ q_rank) {
  pdl_enter();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < st->n_scan;
  const unsigned act = __ballot_sync(0xffffffffu, live);
  if (!live)
    return;
  uint32_t s, r;
  bin_query(q_in_map[i], act, gd, tab, mask, &s, &r);
  q_slot[i] = s;
  q_rank[i] = r;
}

// ---- last kernel of the voxel filter (scan.cuh) ------------------------------------------------------
constexpr int VR_BLOCK = 256, VR_ROWS = 8, VR_TILE = VR_BLOCK * VR_ROWS;
// BIN: the centroids are also transformed into the map frame and binned for the kNN (k_transform + k_query_bin fused in)
template <bool BIN>
__global__ void __launch_bounds__(VR_BLOCK)
    k_vox_reduce(const float4 *__restrict__ cloud, const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, int *d_n,
                 const VoxGridDesc *__restrict__ vgd, float4 *__restrict__ out, unsigned long long *tile_state, unsigned int *tile_ctr,
                 uint32_t epoch, unsigned int *done, int *d_total, int *d_vox_n, int *d_vox_passthrough, Tf32 T,
                 float4 *__restrict__ in_map, const GridDesc *__restrict__ gd, QSlot *__restrict__ tab, uint32_t mask,
                 uint32_t *__restrict__ q_slot, uint32_t *__restrict__ q_rank) {
  __shared__ uint32_t wsum[VR_BLOCK / 32];
  __shared__ uint32_t s_tile, s_prefix, s_tot, s_last, s_last_len;
  __shared__ uint16_t head_off[VR_TILE];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int n = *d_n;
  const int pass = vgd->passthrough;
  const uint32_t active = n > 0 ? (uint32_t)((n + VR_TILE - 1) / VR_TILE) : 1u;
  if (blockIdx.x < active) { // CTA-uniform
    if (pass) {
      // PCL refuses leaf sizes that overflow its int32 voxel index and returns the input unchanged
      for (int k = 0; k < VR_ROWS; k++) {
        const int i = blockIdx.x * VR_TILE + k * VR_BLOCK + threadIdx.x;
        const bool live = i < n;
        const unsigned act = __ballot_sync(0xffffffffu, live);
        if (live) {
          const float4 c = cloud[i];
          out[i] = c;
          if (BIN) {
            const float4 qm = tf_apply(T, c);
            in_map[i] = qm;
            uint32_t sl, rk;
            bin_query(qm, act, gd, tab, mask, &sl, &rk);
            q_slot[i] = sl;
            q_rank[i] = rk;
          }
        }
      }
      if (blockIdx.x == 0 && threadIdx.x == 0)
        *d_total = n;
    } else {
      if (threadIdx.x == 0)
        s_tile = lb_ticket(tile_ctr, active);
      __syncthreads();
      const uint32_t tile = s_tile;
      const int wbase = (int)tile * VR_TILE + wid * (VR_TILE / (VR_BLOCK / 32));
      uint32_t key[VR_ROWS], bal[VR_ROWS], cnt = 0;
#pragma unroll
      for (int r = 0; r < VR_ROWS; r++) {
        const int i = wbase + r * 32 + lane;
        key[r] = i < n ? keys[i] : 0u;
      }
#pragma unroll
      for (int r = 0; r < VR_ROWS; r++) {
        const int i = wbase + r * 32 + lane;
        // previous key: the neighbouring lane's, or (lane 0) the last key of the previous row / of the preceding chunk
        uint32_t prev = __shfl_up_sync(0xffffffffu, key[r], 1);
        const uint32_t prev_row_last = r > 0 ? __shfl_sync(0xffffffffu, key[r > 0 ? r - 1 : 0], 31) : 0u;
        if (lane == 0)
          prev = r > 0 ? prev_row_last : (i > 0 && i < n ? keys[i - 1] : 0u);
        const bool head = i < n && (i == 0 || key[r] != prev);
        bal[r] = __ballot_sync(0xffffffffu, head);
        cnt += __popc(bal[r]);
      }
      if (lane == 0)
        wsum[wid] = cnt;
      __syncthreads();
      if (wid == 0) {
        uint32_t tot = lane < VR_BLOCK / 32 ? wsum[lane] : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
          tot += __shfl_xor_sync(0xffffffffu, tot, o);
        const uint32_t excl = lb_exclusive_prefix(tile_state, tile, epoch, tot);
        if (lane == 0) {
          s_prefix = excl;
          s_tot = tot;
          if ((long long)tile * VR_TILE + VR_TILE >= (long long)n)
            *d_total = (int)(excl + tot);
        }
      }
      __syncthreads();
      // the tile's run heads, in order, into shared memory; then one thread per head (strided over the CTA, so a thread
      // that owns several heads in the load layout does not serialise their runs): the run is summed sequentially
      // (fp32, input order inside the voxel: the sort is stable)
      uint32_t lpos = 0;
      for (int w = 0; w < wid; w++)
        lpos += wsum[w];
      const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
      for (int r = 0; r < VR_ROWS; r++) {
        if ((bal[r] >> lane) & 1u)
          head_off[lpos + __popc(bal[r] & lt)] = (uint16_t)(wid * (VR_TILE / (VR_BLOCK / 32)) + r * 32 + lane);
        lpos += __popc(bal[r]);
      }
      __syncthreads();
      const uint32_t tile_heads = s_tot, prefix = s_prefix;
      const int tile_base = (int)tile * VR_TILE;
      // length of the tile's last run (it may continue into the following tiles): found by one warp, 32 keys per round trip
      if (wid == VR_BLOCK / 32 - 1 && tile_heads > 0) {
        const int i = tile_base + (int)head_off[tile_heads - 1];
        const uint32_t kk = keys[i];
        int len = 0;
        for (int j = i;; j += 32) {
          const unsigned m = __ballot_sync(0xffffffffu, j + lane < n && keys[j + lane] == kk);
          len += (m == 0xffffffffu) ? 32 : __ffs(~m) - 1;
          if (m != 0xffffffffu)
            break;
        }
        if (lane == 0)
          s_last_len = len;
      }
      __syncthreads();
      // one thread per run head (strided over the CTA, so a thread that owns several heads in the load layout does not
      // serialise their runs); the run is added sequentially in fp32 in input order (the sort is stable). Its length is known
      // (next head), so there is no data-dependent branch between the loads: 16 gathers per batch, the indices of the next
      // batch fetched while this one is added (a key comparison per batch put three dependent round trips on every 16 points)
      for (uint32_t kb = 0; kb < tile_heads; kb += VR_BLOCK) {
        const uint32_t k = kb + threadIdx.x;
        const bool live = k < tile_heads;
        const unsigned act = __ballot_sync(0xffffffffu, live);
        if (!live)
          continue;
        const int i = tile_base + (int)head_off[k];
        const int len = k + 1 < tile_heads ? (int)head_off[k + 1] - (int)head_off[k] : (int)s_last_len;
        float sx = 0.f, sy = 0.f, sz = 0.f, sw = 0.f;
        if (len <= 4) { // most voxels of a fine grid (and of an already down-sampled map): 4 gathers instead of 16
          float4 p[4];
#pragma unroll
          for (int u = 0; u < 4; u++)
            p[u] = cloud[vals[i + min(u, len - 1)]]; // clamped: always a valid load
#pragma unroll
          for (int u = 0; u < 4; u++) {
            const bool use = u < len;
            const float ax = __fadd_rn(sx, p[u].x), ay = __fadd_rn(sy, p[u].y), az = __fadd_rn(sz, p[u].z), aw = __fadd_rn(sw, p[u].w);
            sx = use ? ax : sx, sy = use ? ay : sy, sz = use ? az : sz, sw = use ? aw : sw;
          }
        } else {
          uint32_t v[16];
#pragma unroll
          for (int u = 0; u < 16; u++)
            v[u] = vals[i + min(u, len - 1)];
          for (int j0 = 0; j0 < len; j0 += 16) {
            float4 p[16];
#pragma unroll
            for (int u = 0; u < 16; u++)
              p[u] = cloud[v[u]];
            if (j0 + 16 < len) {
#pragma unroll
              for (int u = 0; u < 16; u++)
                v[u] = vals[i + min(j0 + 16 + u, len - 1)];
            }
#pragma unroll
            for (int u = 0; u < 16; u++) {
              const bool use = j0 + u < len;
              const float ax = __fadd_rn(sx, p[u].x), ay = __fadd_rn(sy, p[u].y), az = __fadd_rn(sz, p[u].z), aw = __fadd_rn(sw, p[u].w);
              sx = use ? ax : sx, sy = use ? ay : sy, sz = use ? az : sz, sw = use ? aw : sw;
            }
          }
        }
        const float c = (float)len;
        const float4 ctr = make_float4(__fdiv_rn(sx, c), __fdiv_rn(sy, c), __fdiv_rn(sz, c), __fdiv_rn(sw, c));
        out[prefix + k] = ctr;
        if (BIN) {
          const float4 qm = tf_apply(T, ctr);
          in_map[prefix + k] = qm;
          uint32_t sl, rk;
          bin_query(qm, act, gd, tab, mask, &sl, &rk);
          q_slot[prefix + k] = sl;
          q_rank[prefix + k] = rk;
        }
      }
    }
  }
  // last CTA out publishes the new count (every CTA has read the old one by then)
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0)
    s_last = atomicAdd(done, 1u) == gridDim.x - 1u;
  __syncthreads();
  if (!s_last || threadIdx.x != 0)
    return;
  __threadfence();
  *done = 0u;
  const int total = *(volatile int *)d_total;
  if (d_vox_n)
    *d_vox_n = pass ? 0 : total;
  if (d_vox_passthrough)
    *d_vox_passthrough = pass;
  *d_n = total;
}


// ------------------------------------------------------------------------------------------------
// One-pass front end of the un-down-sampled path: LiDARData's range filter (LidarTypes.cpp:36-55) as a stable compaction
// with decoupled look-back (tile b publishes its count, then the running prefix; a later tile adds counts backwards until
// it meets a prefix), the fp32 transform into the map frame (LidarHelper.cpp:192) and the query binning, all on the
// registers that hold the freshly loaded points. Replaces k_range_count + k_range_compact + k_query_bin (3 launches, 2 extra
// passes over the scan). BIN = false stops after the compaction (a down-sampling pass follows, or there is no map yet). Tiles are handed out by an atomic ticket, so every tile a CTA waits for belongs to a CTA that is
// already running: the look-back cannot deadlock whatever the grid size. States carry a per-scan epoch: no reset pass.
// ------------------------------------------------------------------------------------------------
template <bool VEC4, bool BIN>
__global__ void __launch_bounds__(RF_BLOCK)
    k_range_filter_bin(const float *__restrict__ raw, int stride, int n, float t_rel, double rmin, double rmax, ScanState *__restrict__ st,
                       float4 *__restrict__ cloud, float4 *__restrict__ orig, Tf32 T, float4 *__restrict__ in_map,
                       unsigned long long *tile_state, unsigned int *tile_ctr, uint32_t epoch,
                       const GridDesc *__restrict__ gd, QSlot *__restrict__ tab, uint32_t mask, uint32_t *__restrict__ q_slot,
                       uint32_t *__restrict__ q_rank, float vg_leaf, uint32_t *vg_bbox, unsigned int *vg_done,
                       VoxGridDesc *__restrict__ vg) {
  __shared__ uint32_t wsum[RF_BLOCK / 32];
  __shared__ uint32_t s_tile, s_prefix;
  pdl_enter();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0)
    s_tile = lb_ticket(tile_ctr, gridDim.x);
  __syncthreads();
  const uint32_t tile = s_tile;
  const int wbase = (int)tile * RF_TILE + wid * (RF_TILE / (RF_BLOCK / 32));
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
  if (wid == 0) {
    uint32_t tot = lane < RF_BLOCK / 32 ? wsum[lane] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      tot += __shfl_xor_sync(0xffffffffu, tot, o);
    const uint32_t excl = lb_exclusive_prefix(tile_state, tile, epoch, tot);
    if (lane == 0) {
      s_prefix = excl;
      if (tile == 0) {
        st->n_raw = n;
        st->vox_n = 0;
        st->vox_passthrough = 0;
      }
      if (tile == gridDim.x - 1) {
        st->n_orig = (int)(excl + tot);
        st->n_scan = (int)(excl + tot);
      }
    }
  }
  __syncthreads();
  uint32_t pos = s_prefix;
  for (int w = 0; w < wid; w++)
    pos += wsum[w];
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < 8; r++) {
    if ((bal[r] >> lane) & 1u) {
      const uint32_t o = pos + __popc(bal[r] & lt);
      const float4 q = make_float4(pt[r].x, pt[r].y, pt[r].z, t_rel);
      cloud[o] = q;
      orig[o] = q;
      if (BIN) {
        const float4 qm = tf_apply(T, q);
        in_map[o] = qm;
        uint32_t s, rk;
        bin_query(qm, bal[r], gd, tab, mask, &s, &rk);
        q_slot[o] = s;
        q_rank[o] = rk;
      } else if (in_map) { // transform known but no binning yet (no map, or a down-sampling pass follows)
        in_map[o] = tf_apply(T, q);
      }
    }
    pos += __popc(bal[r]);
  }
  if (!BIN && vg_leaf > 0.f) {
    // a voxel filter follows (one-call path with down-sampling): its first kernel (k_bbox_voxgrid: bounds of the kept points,
    // voxel grid by the CTA that finishes last) is done here on the registers that hold the points
    __shared__ float smn[RF_BLOCK / 32][3], smx[RF_BLOCK / 32][3];
    __shared__ uint32_t s_last;
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int r = 0; r < 8; r++) {
      if ((bal[r] >> lane) & 1u) {
        mn[0] = fminf(mn[0], pt[r].x), mx[0] = fmaxf(mx[0], pt[r].x);
        mn[1] = fminf(mn[1], pt[r].y), mx[1] = fmaxf(mx[1], pt[r].y);
        mn[2] = fminf(mn[2], pt[r].z), mx[2] = fmaxf(mx[2], pt[r].z);
      }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
        mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
      }
    }
    if (lane == 0) {
#pragma unroll
      for (int a = 0; a < 3; a++)
        smn[wid][a] = mn[a], smx[wid][a] = mx[a];
    }
    __syncthreads();
    if (threadIdx.x < 3) {
      const int a = threadIdx.x;
      float lo = INFINITY, hi = -INFINITY;
      for (int w = 0; w < RF_BLOCK / 32; w++)
        lo = fminf(lo, smn[w][a]), hi = fmaxf(hi, smx[w][a]);
      if (lo <= hi) {
        atomicMin(&vg_bbox[a], f2ord(lo));
        atomicMax(&vg_bbox[3 + a], f2ord(hi));
      }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
      s_last = atomicAdd(vg_done, 1u) == gridDim.x - 1u;
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
      __threadfence();
      uint32_t bb[6];
      for (int c = 0; c < 6; c++) {
        bb[c] = *(volatile uint32_t *)&vg_bbox[c];
        vg_bbox[c] = c < 3 ? 0xffffffffu : 0u;
      }
      voxgrid_setup_dev(bb, *(volatile int *)&st->n_scan, vg_leaf, vg);
      *vg_done = 0u;
    }
  }
}

// every occupied slot gets a contiguous range of the sorted query array and ceil(cnt/32) warp units;
// one atomic pair per block (block-level aggregation)
__global__ void __launch_bounds__(256) k_unit_alloc(QSlot *__restrict__ tab, int nslots, ScanState *__restrict__ st, int ph,
                                                   KnnUnit *__restrict__ units, int units_cap) {
  pdl_enter();
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t cnt = (s < nslots) ? tab[s].cnt : 0u;
  uint32_t nu = (cnt + 31u) >> 5;
  uint32_t tq, tu;
  uint32_t eq = block_excl_scan<256>(cnt, &tq);
  uint32_t eu = block_excl_scan<256>(nu, &tu);
  __shared__ uint32_t bq, bu;
  if (threadIdx.x == 0) {
    bq = tq ? (uint32_t)atomicAdd(&st->n_qsorted[ph], (int)tq) : 0u;
    bu = tu ? (uint32_t)atomicAdd(&st->n_units[ph], (int)tu) : 0u;
  }
  __syncthreads();
  if (cnt) {
    uint32_t qb = bq + eq, ub = bu + eu;
    tab[s].qbase = qb;
    int cx, cy, cz;
    cell_unkey(tab[s].key, &cx, &cy, &cz);
    for (uint32_t j = 0; j < nu; j++) {
      if ((int)(ub + j) < units_cap) {
        KnnUnit u;
        u.qbase = (int)(qb + 32u * j);
        u.cnt = (int)min(32u, cnt - 32u * j);
        u.cx = cx, u.cy = cy, u.cz = cz;
        units[ub + j] = u;
      }
    }
  }
}

__global__ void __launch_bounds__(256) k_query_scatter(const float4 *__restrict__ q_in_map, const ScanState *__restrict__ st,
                                                      const QSlot *__restrict__ tab, const uint32_t *__restrict__ q_slot,
                                                      const uint32_t *__restrict__ q_rank, float4 *__restrict__ q_sorted) {
  pdl_enter();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= st->n_scan)
    return;
  float4 q = q_in_map[i];
  q.w = __int_as_float(i);
  q_sorted[tab[q_slot[i]].qbase + q_rank[i]] = q;
}

// ------------------------------------------------------------------------------------------------
// Exact k-NN — replaces KD_TREE::Nearest_Search (ikd_Tree.cpp:358-389, :890-1032) as called by
// LidarHelper::get_neighbors (LidarHelper.cpp:268-289).
// One warp per unit (<= 32 queries of one grid cell). The warp walks cubes of cells of growing half-width
// s = 1..R around the unit's cell; every candidate is loaded once per warp from a warp-uniform address
// (one broadcast wavefront) and tested by all lanes. After level s every map point NOT yet visited is at
// least g_s = (fmin + s) * cell away from the lane's query (fmin = distance of the query to the nearest face of
// its own cell, in cells), so a lane whose k-th best satisfies d2_k < g_s^2 (1 - 1e-6) is exact and complete;
// the warp stops when all its lanes are. R covers sqrt(max_d2): queries still incomplete after level R have
// d2_k > max_d2 and are rejected by get_neighbors anyway (LidarHelper.cpp:284).
// Order: (d2, map index) ascending — total, so results do not depend on the visiting order.
// ------------------------------------------------------------------------------------------------
template <int K> struct TopK {
  float d[K];
  int i[K];
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int j = 0; j < K; j++) {
      d[j] = INFINITY;
      i[j] = 0x7fffffff;
    }
  }
  // (d2, index) ascending as ONE unsigned 64-bit compare: squared distances are non-negative, so their bit patterns order
  // like their values (+inf = 0x7f800000 included) and the pair (bits(d2), index) orders like the 64-bit number
  // bits(d2) << 32 | index. The compiler keeps d and i in separate registers and emits a two-instruction wide compare
  // instead of FSETP.LT + FSETP.EQ + ISETP.LT + a predicate combine.
  static __device__ __forceinline__ unsigned long long key(float d2, int idx) {
    return ((unsigned long long)__float_as_uint(d2) << 32) | (unsigned long long)(unsigned int)idx;
  }
  __device__ __forceinline__ void offer(float d2, int idx) {
    if (key(d2, idx) < key(d[K - 1], i[K - 1])) {
      d[K - 1] = d2;
      i[K - 1] = idx;
#pragma unroll
      for (int s = K - 1; s > 0; --s) {
        const bool sw = key(d[s], i[s]) < key(d[s - 1], i[s - 1]);
        float td = sw ? d[s - 1] : d[s];
        int ti = sw ? i[s - 1] : i[s];
        d[s - 1] = sw ? d[s] : d[s - 1];
        i[s - 1] = sw ? i[s] : i[s - 1];
        d[s] = td;
        i[s] = ti;
      }
    }
  }
};

constexpr int KNN_WARPS = 4;

// The cube walk of one unit (all 32 lanes of a warp call it together). `q` = this lane's query, `active` = lane owns one.
// R = last level to visit (ceil(sqrt(max_d2) / cell)). Leaves the lane's k best in `best`.
template <int K>
__device__ __forceinline__ void knn_walk(const KnnUnit &un, const float4 q, const bool active, const GridDesc &g, const int R,
                                         const uint32_t *__restrict__ cell_start, const float4 *__restrict__ cell_pts, TopK<K> &best) {
  const int lane = threadIdx.x & 31;
  // distance (in cells) from the query to the nearest face of its own cell
  double fx = ((double)q.x - g.ox) * g.inv_cell - (double)un.cx;
  double fy = ((double)q.y - g.oy) * g.inv_cell - (double)un.cy;
  double fz = ((double)q.z - g.oz) * g.inv_cell - (double)un.cz;
  double fmn = fmin(fmin(fmin(fx, 1.0 - fx), fmin(fy, 1.0 - fy)), fmin(fz, 1.0 - fz));
  fmn = fmax(0.0, fmn);
  best.init();

  // levels below s_lo cannot touch the grid (Chebyshev distance from the unit's cell to the grid box); at s_hi the
  // cube covers the whole grid, so every map point has been visited and the search is exact whatever max_d2 is
  const int ox_ = max(max(-un.cx, un.cx - (g.dimx - 1)), 0), oy_ = max(max(-un.cy, un.cy - (g.dimy - 1)), 0),
            oz_ = max(max(-un.cz, un.cz - (g.dimz - 1)), 0);
  const int s_lo = max(1, max(ox_, max(oy_, oz_)));
  const int s_hi = min(R, max(max(max(un.cx, g.dimx - 1 - un.cx), max(un.cy, g.dimy - 1 - un.cy)), max(un.cz, g.dimz - 1 - un.cz)));
  // level 0: the unit's own cell first. Its points are the likeliest neighbours of all 32 queries, so the k-th best of
  // every lane is tight before the 26 surrounding cells are scanned and most of their candidates fail the first compare
  // of offer() for the whole warp (the insertion network then does not issue at all).
  const bool own_in_grid = un.cx >= 0 && un.cx < g.dimx && un.cy >= 0 && un.cy < g.dimy && un.cz >= 0 && un.cz < g.dimz;
  if (own_in_grid) {
    const int c0 = (un.cz * g.dimy + un.cy) * g.dimx + un.cx;
    const uint32_t beg = cell_start[c0], end = cell_start[c0 + 1];
    for (uint32_t j = beg; j < end; j++) {
      const float4 c = __ldg(&cell_pts[j]);
      best.offer(dist2_ref(c.x, c.y, c.z, q.x, q.y, q.z), __float_as_int(c.w));
    }
  }
  for (int s = s_lo; s <= s_hi; s++) {
    // rows (y, z) of the cube clipped to the grid
    const int y0 = max(un.cy - s, 0), y1 = min(un.cy + s, g.dimy - 1);
    const int z0 = max(un.cz - s, 0), z1 = min(un.cz + s, g.dimz - 1);
    const int nz = z1 - z0 + 1;
    const int nrows = (y1 >= y0 && nz > 0) ? (y1 - y0 + 1) * nz : 0;
    for (int rb = 0; rb < nrows; rb += 32) {
      // lane-parallel row set-up: row t -> (y, z); segment A and (for inner rows of levels > s_lo) segment B
      const int t = rb + lane;
      uint32_t begA = 0, endA = 0, begB = 0, endB = 0;
      if (t < nrows) {
        const int y = y0 + t / nz, z = z0 + (t - (t / nz) * nz);
        const int dy = y - un.cy, dz = z - un.cz;
        const int rowbase = (z * g.dimy + y) * g.dimx;
        const bool outer = (s == s_lo) || (abs(dy) == s) || (abs(dz) == s);
        if (outer && own_in_grid && dy == 0 && dz == 0) {
          // the row through the unit's own cell (only at the first level): skip the cell done at level 0
          const int x0 = max(un.cx - s, 0), x1 = min(un.cx + s, g.dimx - 1);
          begA = cell_start[rowbase + x0];
          endA = cell_start[rowbase + un.cx];
          begB = cell_start[rowbase + un.cx + 1];
          endB = cell_start[rowbase + x1 + 1];
        } else if (outer) {
          int x0 = max(un.cx - s, 0), x1 = min(un.cx + s, g.dimx - 1);
          if (x0 <= x1) {
            begA = cell_start[rowbase + x0];
            endA = cell_start[rowbase + x1 + 1];
          }
        } else {
          int xa = un.cx - s, xb = un.cx + s;
          if (xa >= 0 && xa < g.dimx) {
            begA = cell_start[rowbase + xa];
            endA = cell_start[rowbase + xa + 1];
          }
          if (xb >= 0 && xb < g.dimx) {
            begB = cell_start[rowbase + xb];
            endB = cell_start[rowbase + xb + 1];
          }
        }
      }
      // which lanes hold non-empty segments
      unsigned mA = __ballot_sync(0xffffffffu, endA > begA);
      unsigned mB = __ballot_sync(0xffffffffu, endB > begB);
      for (int pass = 0; pass < 2; pass++) {
        unsigned m = pass ? mB : mA;
        while (m) {
          const int src = __ffs(m) - 1;
          m &= m - 1;
          const uint32_t beg = __shfl_sync(0xffffffffu, pass ? begB : begA, src);
          const uint32_t end = __shfl_sync(0xffffffffu, pass ? endB : endA, src);
          uint32_t j = beg;
          for (; j + 4 <= end; j += 4) {
            const float4 c0 = __ldg(&cell_pts[j]), c1 = __ldg(&cell_pts[j + 1]), c2 = __ldg(&cell_pts[j + 2]),
                         c3 = __ldg(&cell_pts[j + 3]);
            best.offer(dist2_ref(c0.x, c0.y, c0.z, q.x, q.y, q.z), __float_as_int(c0.w));
            best.offer(dist2_ref(c1.x, c1.y, c1.z, q.x, q.y, q.z), __float_as_int(c1.w));
            best.offer(dist2_ref(c2.x, c2.y, c2.z, q.x, q.y, q.z), __float_as_int(c2.w));
            best.offer(dist2_ref(c3.x, c3.y, c3.z, q.x, q.y, q.z), __float_as_int(c3.w));
          }
          for (; j < end; j++) {
            const float4 c0 = __ldg(&cell_pts[j]);
            best.offer(dist2_ref(c0.x, c0.y, c0.z, q.x, q.y, q.z), __float_as_int(c0.w));
          }
        }
      }
    }
    // completeness test for this level
    const double gs = (fmn + (double)s) * g.cell;
    const bool done = !active || ((double)best.d[K - 1] < gs * gs * (1.0 - 1e-6));
    if (__all_sync(0xffffffffu, done))
      break;
  }
}

// ---- TMA 1-D bulk copy global -> shared, completion on an mbarrier (cp.async.bulk; SASS UBLKCP). One thread issues the copy
// of a whole 32 KB tile; nobody's registers, load queue or memory fences are involved while it is in flight - which is the
// point here: a __threadfence() waits for the issuing thread's outstanding LOADS too, so a register prefetch of the next tile
// over PCIe (11 us) stalled the publication of the current one. ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, unsigned long long *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
               "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, uint32_t parity) {
  asm volatile("{\n"
               ".reg .pred p;\n"
               "MBAR_WAIT:\n"
               "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
               "@p bra MBAR_DONE;\n"
               "bra MBAR_WAIT;\n"
               "MBAR_DONE:\n"
               "}" ::"r"(smem_u32(bar)),
               "r"(parity)
               : "memory");
}


// ------------------------------------------------------------------------------------------------
// The same cube walk with the candidates STAGED IN SHARED MEMORY by TMA bulk copies (north_star's design for the search).
// Why: knn_walk loads every candidate from a warp-uniform global address, so a warp has at most 4 x 16 bytes in flight and a
// unit's walk is a chain of ~45 dependent L2 round trips (9 rows x ~5 groups of 4): at one unit per warp the search phase of a
// 131 072-ray scan lasted 25-30 us although its instructions need 13 us of issue (profiles/r2_ncu.md). Here the lanes compute
// the cube's rows (x-runs of cells = contiguous ranges of cell_pts), every lane hands its range(s) to the copy engine
// (cp.async.bulk -> the warp's own buffer, completion on the warp's mbarrier), and the warp then tests the flat candidate list
// from shared memory (one broadcast LDS.128 per candidate): one round trip per level instead of ~45.
// The buffer is two halves of KW_CH candidates with one mbarrier each: chunk c + 2 is requested as soon as chunk c has been
// tested. Arithmetic, completeness rule and (d2, index) order are those of knn_walk: results are bit-identical.
// The unit's own cell is no special case here: at the first level the whole cube is staged, the row through the own cell first
// (it holds the likeliest neighbours, so the k-th best is tight when the other rows are tested).
// ------------------------------------------------------------------------------------------------
constexpr int KW_CH = 192;                                     // candidates per half buffer
constexpr size_t KW_WARP_BYTES = 2 * KW_CH * sizeof(float4);   // 6 KB per warp

template <int K>
__device__ __forceinline__ void knn_walk_staged(const KnnUnit &un, const float4 q, const bool active, const GridDesc &g, const int R,
                                                const uint32_t *__restrict__ cell_start, const float4 *__restrict__ cell_pts,
                                                TopK<K> &best, float4 *buf, unsigned long long *bar, uint32_t &par,
                                                unsigned long long *prof = nullptr) {
  const int lane = threadIdx.x & 31;
#if defined(ML_KNN_PROF) && ML_KNN_PROF
#define KW_T(i)                                                                                                                       \
  {                                                                                                                                   \
    unsigned long long t_;                                                                                                            \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                                                            \
    if (prof)                                                                                                                         \
      prof[i] += t_ - kw_t;                                                                                                           \
    kw_t = t_;                                                                                                                        \
  }
  unsigned long long kw_t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(kw_t));
#else
#define KW_T(i)
#endif
  double fx = ((double)q.x - g.ox) * g.inv_cell - (double)un.cx;
  double fy = ((double)q.y - g.oy) * g.inv_cell - (double)un.cy;
  double fz = ((double)q.z - g.oz) * g.inv_cell - (double)un.cz;
  double fmn = fmin(fmin(fmin(fx, 1.0 - fx), fmin(fy, 1.0 - fy)), fmin(fz, 1.0 - fz));
  fmn = fmax(0.0, fmn);
  best.init();
  const int ox_ = max(max(-un.cx, un.cx - (g.dimx - 1)), 0), oy_ = max(max(-un.cy, un.cy - (g.dimy - 1)), 0),
            oz_ = max(max(-un.cz, un.cz - (g.dimz - 1)), 0);
  const int s_lo = max(1, max(ox_, max(oy_, oz_)));
  const int s_hi = min(R, max(max(max(un.cx, g.dimx - 1 - un.cx), max(un.cy, g.dimy - 1 - un.cy)), max(un.cz, g.dimz - 1 - un.cz)));
  const bool own_in_grid = un.cx >= 0 && un.cx < g.dimx && un.cy >= 0 && un.cy < g.dimy && un.cz >= 0 && un.cz < g.dimz;
  for (int s = s_lo; s <= s_hi; s++) {
    const int y0 = max(un.cy - s, 0), y1 = min(un.cy + s, g.dimy - 1);
    const int z0 = max(un.cz - s, 0), z1 = min(un.cz + s, g.dimz - 1);
    const int nz = z1 - z0 + 1;
    const int nrows = (y1 >= y0 && nz > 0) ? (y1 - y0 + 1) * nz : 0;
    // own_in_grid implies s_lo = 1: at most 9 rows, the own row's index fits the first batch
    const int t_own = (s == s_lo && own_in_grid) ? (un.cy - y0) * nz + (un.cz - z0) : 0;
    for (int rb = 0; rb < nrows; rb += 32) {
      int t = rb + lane;
      if (rb == 0)
        t = (t == 0) ? t_own : (t == t_own ? 0 : t); // the own row goes first
      uint32_t begA = 0, lenA = 0, begB = 0, lenB = 0;
      if (t < nrows) {
        const int y = y0 + t / nz, z = z0 + (t - (t / nz) * nz);
        const int dy = y - un.cy, dz = z - un.cz;
        const int rowbase = (z * g.dimy + y) * g.dimx;
        const bool outer = (s == s_lo) || (abs(dy) == s) || (abs(dz) == s);
        if (outer) {
          const int x0 = max(un.cx - s, 0), x1 = min(un.cx + s, g.dimx - 1);
          if (x0 <= x1) {
            begA = cell_start[rowbase + x0];
            lenA = cell_start[rowbase + x1 + 1] - begA;
          }
        } else {
          const int xa = un.cx - s, xb = un.cx + s;
          if (xa >= 0 && xa < g.dimx) {
            begA = cell_start[rowbase + xa];
            lenA = cell_start[rowbase + xa + 1] - begA;
          }
          if (xb >= 0 && xb < g.dimx) {
            begB = cell_start[rowbase + xb];
            lenB = cell_start[rowbase + xb + 1] - begB;
          }
        }
      }
      // flat candidate list of the batch: lane order, A before B
      uint32_t incl = lenA + lenB;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o)
          incl += v;
      }
      const uint32_t T = __shfl_sync(0xffffffffu, incl, 31);
      KW_T(0); // rows set up: cell_start loads + scan
      const uint32_t offA = incl - lenA - lenB, offB = offA + lenA;
      const int nch = (int)((T + KW_CH - 1) / KW_CH);
      auto issue = [&](int c) {
        const int h = c & 1;
        const uint32_t lo = (uint32_t)c * KW_CH, hi = min(T, lo + (uint32_t)KW_CH);
        if (lane == 0)
          mbar_expect_tx(&bar[h], (hi - lo) * (uint32_t)sizeof(float4));
        __syncwarp();
        {
          const uint32_t a = max(offA, lo), b = min(offA + lenA, hi);
          if (a < b)
            bulk_g2s(buf + h * KW_CH + (a - lo), cell_pts + begA + (a - offA), (b - a) * (uint32_t)sizeof(float4), &bar[h]);
        }
        {
          const uint32_t a = max(offB, lo), b = min(offB + lenB, hi);
          if (a < b)
            bulk_g2s(buf + h * KW_CH + (a - lo), cell_pts + begB + (a - offB), (b - a) * (uint32_t)sizeof(float4), &bar[h]);
        }
      };
      if (nch > 0)
        issue(0);
      if (nch > 1)
        issue(1);
      for (int c = 0; c < nch; c++) {
        const int h = c & 1;
        mbar_wait(&bar[h], (par >> h) & 1u);
        par ^= 1u << h;
        KW_T(1); // candidates staged
        const int cnt = (int)min(T - (uint32_t)c * KW_CH, (uint32_t)KW_CH);
        const float4 *p = buf + h * KW_CH;
        int j = 0;
        for (; j + 4 <= cnt; j += 4) {
          const float4 c0 = p[j], c1 = p[j + 1], c2 = p[j + 2], c3 = p[j + 3];
          best.offer(dist2_ref(c0.x, c0.y, c0.z, q.x, q.y, q.z), __float_as_int(c0.w));
          best.offer(dist2_ref(c1.x, c1.y, c1.z, q.x, q.y, q.z), __float_as_int(c1.w));
          best.offer(dist2_ref(c2.x, c2.y, c2.z, q.x, q.y, q.z), __float_as_int(c2.w));
          best.offer(dist2_ref(c3.x, c3.y, c3.z, q.x, q.y, q.z), __float_as_int(c3.w));
        }
        for (; j < cnt; j++) {
          const float4 c0 = p[j];
          best.offer(dist2_ref(c0.x, c0.y, c0.z, q.x, q.y, q.z), __float_as_int(c0.w));
        }
        __syncwarp(); // every lane has read the half before it is handed back to the copy engine
        KW_T(2); // candidates tested
        if (c + 2 < nch)
          issue(c + 2);
      }
    }
    const double gs = (fmn + (double)s) * g.cell;
    const bool done = !active || ((double)best.d[K - 1] < gs * gs * (1.0 - 1e-6));
    if (__all_sync(0xffffffffu, done))
      break;
  }
}

// ------------------------------------------------------------------------------------------------
// The cube walk of ONE query by ONE thread (no grouping, no shared memory): every lane of a warp walks the cube of its own
// query's cell. Same levels, same rows, same completeness rule, same arithmetic and (d2, index) order as knn_walk: identical
// results. Lanes whose queries share a cell run in lock step and their loads coalesce into one broadcast; lanes in different
// cells differ in trip counts only. The ranges of the next row are requested before the candidates of the current row are
// tested, so a level costs (rows + 1) dependent round trips.
// Used by k_scan_tile (tile.cuh), where a warp's 32 queries are 32 consecutive rays: a warp-per-cell unit of the grouped walk
// holds 15 queries and 26 candidates on average on the 64-beam workload, and its cost was all latency and queueing
// (profiles/r2_ncu.md).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
template <int K, bool PREFETCH = false>
__device__ __forceinline__ void knn_walk_lane(const float4 q, const int cx, const int cy, const int cz, const GridDesc &g, const int R,
                                              const uint32_t *__restrict__ cell_start, const float4 *__restrict__ cell_pts,
                                              TopK<K> &best) {
  double fx = ((double)q.x - g.ox) * g.inv_cell - (double)cx;
  double fy = ((double)q.y - g.oy) * g.inv_cell - (double)cy;
  double fz = ((double)q.z - g.oz) * g.inv_cell - (double)cz;
  double fmn = fmin(fmin(fmin(fx, 1.0 - fx), fmin(fy, 1.0 - fy)), fmin(fz, 1.0 - fz));
  fmn = fmax(0.0, fmn);
  best.init();
  const int ox_ = max(max(-cx, cx - (g.dimx - 1)), 0), oy_ = max(max(-cy, cy - (g.dimy - 1)), 0), oz_ = max(max(-cz, cz - (g.dimz - 1)), 0);
  const int s_lo = max(1, max(ox_, max(oy_, oz_)));
  const int s_hi = min(R, max(max(max(cx, g.dimx - 1 - cx), max(cy, g.dimy - 1 - cy)), max(cz, g.dimz - 1 - cz)));
  for (int s = s_lo; s <= s_hi; s++) {
    const int y0 = max(cy - s, 0), y1 = min(cy + s, g.dimy - 1);
    const int z0 = max(cz - s, 0), z1 = min(cz + s, g.dimz - 1);
    const int x0 = max(cx - s, 0), x1 = min(cx + s, g.dimx - 1);
    const int xa = cx - s, xb = cx + s;
    // row (y, z) -> up to two ranges of cell_pts: the whole x-run on the faces of the cube (and at the first level), else the
    // two end cells
    auto row_ranges = [&](int y, int z, uint32_t &bA, uint32_t &eA, uint32_t &bB, uint32_t &eB) {
      bA = eA = bB = eB = 0u;
      const int rowbase = (z * g.dimy + y) * g.dimx;
      const bool outer = (s == s_lo) || (abs(y - cy) == s) || (abs(z - cz) == s);
      if (outer) {
        if (x0 <= x1) {
          bA = __ldg(&cell_start[rowbase + x0]);
          eA = __ldg(&cell_start[rowbase + x1 + 1]);
        }
      } else {
        if (xa >= 0 && xa < g.dimx) {
          bA = __ldg(&cell_start[rowbase + xa]);
          eA = __ldg(&cell_start[rowbase + xa + 1]);
        }
        if (xb >= 0 && xb < g.dimx) {
          bB = __ldg(&cell_start[rowbase + xb]);
          eB = __ldg(&cell_start[rowbase + xb + 1]);
        }
      }
    };
    if (y1 >= y0 && z1 >= z0) {
      // Latency: a thread's walk is a chain of dependent loads (ranges of a row, then its candidates), ~2 round trips per row.
      // Two prefetch passes turn that into ~2 round trips per LEVEL: first the cell_start lines of every row (their addresses
      // need no load), then - reading the ranges from L1 - the cell_pts lines of every row; the walk proper then hits L1.
      // Only for the small cubes of the first levels (9 to 25 rows). MEASURED SLOWER and therefore off: on the 64-beam workload
      // the walk took 29 us instead of 16.5 us per tile (device step 73.8 us against 61.7 us) - the prefetches queue behind the
      // same load pipe the walk needs and most of a tile's lines are in L1 already.
      if (PREFETCH && (y1 - y0 + 1) * (z1 - z0 + 1) <= 25) {
        for (int y = y0; y <= y1; y++)
          for (int z = z0; z <= z1; z++) {
            const int rowbase = (z * g.dimy + y) * g.dimx;
            prefetch_l1(&cell_start[rowbase + max(xa, 0)]);
            prefetch_l1(&cell_start[rowbase + min(xb, g.dimx - 1) + 1]);
          }
        for (int y = y0; y <= y1; y++)
          for (int z = z0; z <= z1; z++) {
            uint32_t bA, eA, bB, eB;
            row_ranges(y, z, bA, eA, bB, eB);
            for (uint32_t a = bA & ~7u; a < eA; a += 8u) // 8 float4 per 128-byte line
              prefetch_l1(&cell_pts[a]);
            for (uint32_t a = bB & ~7u; a < eB; a += 8u)
              prefetch_l1(&cell_pts[a]);
          }
      }
      int y = y0, z = z0;
      uint32_t nbA, neA, nbB, neB;
      row_ranges(y, z, nbA, neA, nbB, neB);
      for (;;) {
        const uint32_t bA = nbA, eA = neA, bB = nbB, eB = neB;
        // the next row's ranges are requested before this row's candidates are tested
        bool more = true;
        if (++z > z1) {
          z = z0;
          if (++y > y1)
            more = false;
        }
        if (more)
          row_ranges(y, z, nbA, neA, nbB, neB);
        uint32_t j = bA;
        for (; j + 4 <= eA; j += 4) {
          const float4 c0 = __ldg(&cell_pts[j]), c1 = __ldg(&cell_pts[j + 1]), c2 = __ldg(&cell_pts[j + 2]), c3 = __ldg(&cell_pts[j + 3]);
          best.offer(dist2_ref(c0.x, c0.y, c0.z, q.x, q.y, q.z), __float_as_int(c0.w));
          best.offer(dist2_ref(c1.x, c1.y, c1.z, q.x, q.y, q.z), __float_as_int(c1.w));
          best.offer(dist2_ref(c2.x, c2.y, c2.z, q.x, q.y, q.z), __float_as_int(c2.w));
          best.offer(dist2_ref(c3.x, c3.y, c3.z, q.x, q.y, q.z), __float_as_int(c3.w));
        }
        for (; j < eA; j++) {
          const float4 c0 = __ldg(&cell_pts[j]);
          best.offer(dist2_ref(c0.x, c0.y, c0.z, q.x, q.y, q.z), __float_as_int(c0.w));
        }
        for (j = bB; j < eB; j++) {
          const float4 c0 = __ldg(&cell_pts[j]);
          best.offer(dist2_ref(c0.x, c0.y, c0.z, q.x, q.y, q.z), __float_as_int(c0.w));
        }
        if (!more)
          break;
      }
    }
    const double gs = (fmn + (double)s) * g.cell;
    if ((double)best.d[K - 1] < gs * gs * (1.0 - 1e-6))
      break;
  }
}

// What one query hands on (all outputs optional): the parity export (indices, squared distances), get_neighbors' verdict
// (LidarHelper.cpp:280-285) and - what KD_TREE::Nearest_Search hands to compute_plane - the neighbours' COORDINATES
// (ikd_Tree.cpp:383-387) in scan order (16 K contiguous bytes per query), so that the rows kernel reads them coalesced instead of
// gathering by index. A query get_neighbors rejects is marked with a NaN first coordinate (the rows kernel needs no other flag).
template <int K>
__device__ __forceinline__ void knn_emit(const TopK<K> &best, const int qi, const float max_d2, int32_t *__restrict__ knn_idx,
                                         float *__restrict__ knn_d2, uint8_t *__restrict__ knn_ok, const float4 *__restrict__ map_pts,
                                         float4 *__restrict__ nbr) {
  const bool found = best.i[K - 1] != 0x7fffffff;
  if (knn_idx) {
#pragma unroll
    for (int j = 0; j < K; j++) {
      // candidates beyond the search radius are not reported (same convention as the brute-force oracle)
      const bool in = best.i[j] != 0x7fffffff && best.d[j] <= max_d2;
      knn_idx[(size_t)qi * K + j] = in ? best.i[j] : -1;
      knn_d2[(size_t)qi * K + j] = in ? best.d[j] : INFINITY;
    }
  }
  const bool ok = found && !((double)best.d[K - 1] > (double)max_d2);
  if (knn_ok)
    knn_ok[qi] = ok ? 1 : 0;
  if (nbr) {
    if (ok) {
      float4 p[K];
#pragma unroll
      for (int j = 0; j < K; j++)
        p[j] = __ldg(&map_pts[best.i[j]]);
#pragma unroll
      for (int j = 0; j < K; j++)
        nbr[(size_t)qi * K + j] = p[j];
    } else {
      nbr[(size_t)qi * K] = make_float4(__int_as_float(0x7fc00000), 0.f, 0.f, 0.f);
    }
  }
}

template <int K>
__global__ void __launch_bounds__(KNN_WARPS * 32)
    k_knn(const KnnUnit *__restrict__ units, ScanState *__restrict__ st, int ph, const float4 *__restrict__ q_sorted,
          const GridDesc *__restrict__ gd, const uint32_t *__restrict__ cell_start, const float4 *__restrict__ cell_pts, float max_d2,
          int32_t *__restrict__ knn_idx, float *__restrict__ knn_d2, uint8_t *__restrict__ knn_ok, QSlot *__restrict__ qtab, int q_slots,
          const float4 *__restrict__ map_pts, float4 *__restrict__ nbr, unsigned long long *__restrict__ tl) {
  tl_mark(tl, 0);
  pdl_enter();
  tl_mark(tl, 1);
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int n_units = st->n_units[ph];
  // housekeeping for the next kNN call: the query hash table is dead once the queries are in cell order (nobody reads it
  // in this kernel), and the other parity's work-list counters are idle
  for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < q_slots; s += gridDim.x * blockDim.x) {
    qtab[s].key = 0ull;
    qtab[s].cnt = 0u;
    qtab[s].qbase = 0u;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->n_units[ph ^ 1] = 0;
    st->n_qsorted[ph ^ 1] = 0;
    st->knn_next[ph ^ 1] = 0;
  }
  const GridDesc g = *gd;
  const int R = max(1, (int)ceil(sqrt((double)max_d2) * (1.0 + 1e-5) * g.inv_cell));

  // Units differ widely in cost (candidates in reach, depth of the walk): a warp's first unit is static (no atomic on the way
  // in), every further one is taken from a ticket, so the kernel ends when the work does and not with the unluckiest stride.
  // Results are per query and do not depend on who computes them.
  for (int u = warp; u < n_units;) {
    const KnnUnit un = units[u];
    const bool active = lane < un.cnt;
    const float4 q = q_sorted[un.qbase + (active ? lane : 0)];
    TopK<K> best;
    knn_walk<K>(un, q, active, g, R, cell_start, cell_pts, best);
    if (active)
      knn_emit<K>(best, __float_as_int(q.w), max_d2, knn_idx, knn_d2, knn_ok, map_pts, nbr);
    int nxt = 0;
    if (lane == 0)
      nxt = nwarps + atomicAdd(&st->knn_next[ph], 1);
    u = __shfl_sync(0xffffffffu, nxt, 0);
  }
  tl_mark(tl, 2);
}

} // namespace ml
