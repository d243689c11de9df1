// One persistent kernel for the front of the un-down-sampled scan update: LiDARData's range filter (LidarTypes.cpp:36-55) as a
// stable single-pass compaction, the fp32 transform into the map frame (LidarHelper.cpp:192), the grouping of the queries by grid
// cell into warp units, and the exact k-NN (LidarHelper::get_neighbors / KD_TREE::Nearest_Search) of every unit.
// Replaces k_range_filter_bin + k_unit_alloc + k_query_scatter + k_knn (4 launches, a 4 MB global hash table and two extra
// passes over the scan) on the one-call path.
//
// Why one kernel: with the scan in pinned HOST memory the front end is a PCIe read (2.1 MB at ~47 GB/s = 45 us for a
// 131 072-ray scan) and nothing downstream could start before its last byte had arrived. Here the first `producers` CTAs walk
// the 2048-ray tiles IN ORDER (an atomic ticket), so the bytes of tile t arrive before those of tile t + producers; a finished
// tile publishes its units at once and every other warp of the grid is a consumer that takes published units from a second
// ticket and searches them while later tiles are still crossing the bus. A producer CTA turns consumer when the tiles run out.
//
// Producer, per tile (256 threads x 8 rays):
//   load -> range test -> ballots -> decoupled look-back (prims.cuh) for the tile's offset in the compacted clouds -> write
//   pointcloud / pointcloud_original / pointcloud_in_map in scan order -> group the tile's queries by grid cell in a
//   SHARED-MEMORY hash table (warp-aggregated: one probe per distinct cell per warp) -> exclusive scan over the table slots ->
//   reserve a range of the unit list (one global atomic per tile; the tile's range of the cell-ordered query array is its range
//   of the compacted clouds) -> scatter queries, write unit records -> fence -> stamp the units' tags with the scan's epoch ->
//   only then the three clouds (nobody waits for them) -> tiles_done++.
//   A cell that spans two tiles becomes two units (+14 % units on the 64-beam scan at 2048-ray tiles): the result of a query does
//   not depend on its unit.
// Consumer, per warp: u = unit ticket; wait until tag[u] == epoch (or until all tiles are done and u is past the final count);
//   cube walk of update.cuh (bit-identical arithmetic and order), outputs by original scan index.
// Tags carry the epoch of the scan, so nothing is cleared between scans; the control block is reset by the last CTA to leave.
#pragma once
#include "update.cuh"

namespace ml {

struct FusedCtl {
  // written by the producers (a few atomics per tile), read by waiting consumers
  unsigned int tile_next;    // front tiles handed out
  unsigned int tiles_done;   // front tiles whose units are published
  unsigned int unit_reserve; // units reserved so far; final once tiles_done == ntiles
  unsigned int pad0[29];
  // written by the consumers (one atomic per unit): its own 128-byte line
  unsigned int unit_next;    // unit tickets handed out
  unsigned int ctas_out;     // CTAs that have left
  unsigned int pad1[30];
};

constexpr int FK_BLOCK = 512;             // 16 warps; two CTAs per SM
constexpr int FK_ROWS = RF_TILE / FK_BLOCK; // 4 rays per thread: warp w owns 128 consecutive rays, a row is 32 of them
constexpr int FK_SLOTS = 2 * RF_TILE;     // 4096: shared-memory hash table, load factor <= 1/2
constexpr size_t FK_HASH_BYTES = (size_t)FK_SLOTS * (sizeof(unsigned long long) + 2 * sizeof(uint32_t)); // 64 KB
constexpr size_t FK_STAGE_BYTES = (size_t)RF_TILE * sizeof(float4);                                          // 32 KB: the next tile's rays
constexpr size_t FK_SMEM = FK_HASH_BYTES + FK_STAGE_BYTES;                                                   // 96 KB, two CTAs per SM
#ifndef ML_KNN_STAGED
#define ML_KNN_STAGED 1
#endif
// -DML_KNN_PROF=1 (scripts/knn_warp_profile.py): instead of the per-CTA marks every consumer warp records, in the timeline buffer,
// [ns waiting for tickets and tags, ns in the walk, ns emitting, units, candidates offered, first unit start, last unit end, 0]
#ifndef ML_KNN_PROF
#define ML_KNN_PROF 0
#endif
#if ML_KNN_PROF
#define FK_MARK(tl, s) ((void)0)
#else
#define FK_MARK(tl, s) tl_mark(tl, s)
#endif
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
static_assert((FK_BLOCK / 32) * KW_WARP_BYTES <= FK_SMEM, "the consumers' candidate buffers reuse the producer's shared memory");

__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// the raw scan is read exactly once: L2 loads (no L1 allocation)
template <bool VEC4> __device__ __forceinline__ float4 load_raw_cg(const float *raw, int stride, int i) {
  if (VEC4)
    return __ldcg(reinterpret_cast<const float4 *>(raw) + i);
  const float *p = raw + (size_t)i * stride;
  return make_float4(__ldcg(p), __ldcg(p + 1), __ldcg(p + 2), 0.f);
}

template <int K, bool VEC4>
__global__ void __launch_bounds__(FK_BLOCK, 2)
    k_front_knn(const float *__restrict__ raw, int stride, int n, float t_rel, double rmin, double rmax, ScanState *__restrict__ st,
                float4 *__restrict__ cloud, float4 *__restrict__ orig, Tf32 T, float4 *__restrict__ in_map,
                unsigned long long *tile_state, uint32_t epoch, FusedCtl *ctl, int producers, const GridDesc *__restrict__ gd,
                float4 *__restrict__ q_sorted, KnnUnit *__restrict__ units, uint32_t *__restrict__ units_tag,
                const uint32_t *__restrict__ cell_start, const float4 *__restrict__ cell_pts, float max_d2,
                int32_t *__restrict__ knn_idx, float *__restrict__ knn_d2, uint8_t *__restrict__ knn_ok,
                const float4 *__restrict__ map_pts, float4 *__restrict__ nbr, unsigned long long *__restrict__ tl) {
  extern __shared__ __align__(16) unsigned char fk_smem[];
  unsigned long long *hkey = reinterpret_cast<unsigned long long *>(fk_smem); // [FK_SLOTS] 0 = empty
  uint32_t *hcnt = reinterpret_cast<uint32_t *>(hkey + FK_SLOTS);             // [FK_SLOTS] queries of the cell in this tile
  uint32_t *hpre = hcnt + FK_SLOTS;                                           // [FK_SLOTS] (query base) | (unit base) << 16
  float4 *stage = reinterpret_cast<float4 *>(fk_smem + FK_HASH_BYTES);        // [RF_TILE] the next tile's rays (VEC4 input)
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ __align__(8) unsigned long long s_kbar[FK_BLOCK / 32][2]; // consumers: one mbarrier per half of a warp's candidate buffer
  __shared__ uint32_t wsum[FK_BLOCK / 32];
  __shared__ uint32_t s_next, s_prefix, s_gu, s_tu;
  FK_MARK(tl, 0);
  pdl_enter();
  FK_MARK(tl, 1);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t ntiles = (uint32_t)((n + RF_TILE - 1) / RF_TILE);
  const GridDesc g = *gd;

  // ------------------------------------------------------------------------------------------ producer
  if ((int)blockIdx.x < producers && blockIdx.x < ntiles) {
    // A producer's first tile is its block index (no ticket round trip on the way in); the ticket hands out the rest. The rays
    // of the NEXT tile are requested as soon as the current ones have landed, so the transfer (PCIe, with the scan in pinned
    // host memory) runs beside the grouping of the current tile: a few producers keep the link busy and finish their tiles one
    // after the other, which is what lets the consumers start while most of the scan is still on the bus.
    uint32_t tile = blockIdx.x;
    uint32_t phase = 0;
    if (VEC4 && threadIdx.x == 0) {
      mbar_init(&s_bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    float4 pt[FK_ROWS];
    {
      const int wbase = (int)tile * RF_TILE + wid * (32 * FK_ROWS);
#pragma unroll
      for (int r = 0; r < FK_ROWS; r++) {
        const int i = wbase + r * 32 + lane;
        pt[r] = i < n ? load_raw_cg<VEC4>(raw, stride, i) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    for (;;) {
      if (threadIdx.x == 0)
        s_next = (uint32_t)producers + atomicAdd(&ctl->tile_next, 1u); // the next tile's ticket, early
      // clear the table while the loads are in flight
      for (int s = threadIdx.x; s < FK_SLOTS; s += FK_BLOCK) {
        hkey[s] = 0ull;
        hcnt[s] = 0u;
      }
      const int wbase = (int)tile * RF_TILE + wid * (32 * FK_ROWS);
      uint32_t bal[FK_ROWS];
      uint32_t cnt = 0;
#pragma unroll
      for (int r = 0; r < FK_ROWS; r++) {
        const int i = wbase + r * 32 + lane;
        const bool k = i < n && range_keep(pt[r], rmin, rmax);
        bal[r] = __ballot_sync(0xffffffffu, k);
        cnt += __popc(bal[r]);
      }
      if (lane == 0)
        wsum[wid] = cnt;
      __syncthreads();
      FK_MARK(tl, 4); // loads landed, range test done
      const uint32_t next = s_next;
      if (VEC4 && next < ntiles && threadIdx.x == 0) { // the next tile's rays start their journey now (one bulk copy)
        const uint32_t bytes = (uint32_t)min(RF_TILE, n - (int)next * RF_TILE) * (uint32_t)sizeof(float4);
        mbar_expect_tx(&s_bar, bytes);
        bulk_g2s(stage, reinterpret_cast<const float4 *>(raw) + (size_t)next * RF_TILE, bytes, &s_bar);
      }
      // warp 0: the tile's offset in the compacted clouds (decoupled look-back). The other warps go on to the hash fill, which
      // needs no offset: the look-back's L2 round trips hide behind it.
      if (wid == 0) {
        uint32_t tot = lane < FK_BLOCK / 32 ? wsum[lane] : 0u;
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
          if (tile == ntiles - 1) {
            st->n_orig = (int)(excl + tot);
            st->n_scan = (int)(excl + tot);
          }
        }
      }
      // hash fill: cell of every kept query (its map-frame position is recomputed for the stores below: cheaper than keeping it
      // live), warp-aggregated insert; sr = slot | rank << 16. In phases over the rows, so that the long-latency steps of
      // different rows overlap: keys, peer sets, slot claims, counts, broadcast.
      const uint32_t lt = (1u << lane) - 1u;
      unsigned long long key[FK_ROWS];
      uint32_t peers[FK_ROWS], sr[FK_ROWS], base[FK_ROWS];
#pragma unroll
      for (int r = 0; r < FK_ROWS; r++) {
        const float4 qm = tf_apply(T, pt[r]);
        const int lim = 1 << 19;
        const int cx = max(-lim, min(lim, cell_coord((double)qm.x, g.ox, g.inv_cell)));
        const int cy = max(-lim, min(lim, cell_coord((double)qm.y, g.oy, g.inv_cell)));
        const int cz = max(-lim, min(lim, cell_coord((double)qm.z, g.oz, g.inv_cell)));
        key[r] = cell_key(cx, cy, cz);
      }
#pragma unroll
      for (int r = 0; r < FK_ROWS; r++) {
        const bool kept = (bal[r] >> lane) & 1u;
        // rays the range gate dropped get a key of their own (a lane number is no cell key): they match nobody
        peers[r] = __match_any_sync(0xffffffffu, kept ? key[r] : (unsigned long long)lane) & bal[r];
      }
#pragma unroll
      for (int r = 0; r < FK_ROWS; r++) { // leaders claim the slot of their cell
        sr[r] = 0u;
        const bool kept = (bal[r] >> lane) & 1u;
        if (kept && lane == __ffs(peers[r]) - 1) {
          uint32_t s = hash64(key[r]) & (FK_SLOTS - 1);
          for (;;) {
            const unsigned long long prev = atomicCAS(&hkey[s], 0ull, key[r]);
            if (prev == 0ull || prev == key[r])
              break;
            s = (s + 1) & (FK_SLOTS - 1);
          }
          sr[r] = s;
        }
      }
#pragma unroll
      for (int r = 0; r < FK_ROWS; r++) { // ... and reserve the ranks of all their peers with one atomicAdd
        const bool kept = (bal[r] >> lane) & 1u;
        base[r] = (kept && lane == __ffs(peers[r]) - 1) ? atomicAdd(&hcnt[sr[r]], (uint32_t)__popc(peers[r])) : 0u;
      }
#pragma unroll
      for (int r = 0; r < FK_ROWS; r++) {
        const bool kept = (bal[r] >> lane) & 1u;
        const int leader = kept ? __ffs(peers[r]) - 1 : lane;
        const uint32_t s = __shfl_sync(0xffffffffu, sr[r], leader);
        const uint32_t b = __shfl_sync(0xffffffffu, base[r], leader);
        sr[r] = kept ? (s | ((b + (uint32_t)__popc(peers[r] & lt)) << 16)) : 0u;
      }
      __syncthreads();
      FK_MARK(tl, 5); // look-back done, hash filled
      // exclusive scan over the slots: queries and units (ceil(cnt / 32)) of every occupied slot, 8 consecutive slots per thread
      uint32_t my_gu = 0, my_tu = 0;
      {
        constexpr int SPT = FK_SLOTS / FK_BLOCK; // 8
        uint32_t lq = 0, lu = 0;
#pragma unroll
        for (int j = 0; j < SPT; j++) {
          const uint32_t c = hcnt[threadIdx.x * SPT + j];
          lq += c;
          lu += (c + 31u) >> 5;
        }
        // one block scan for both: the sums fit 16 bits each (<= 2048)
        uint32_t tot2;
        const uint32_t e2 = block_excl_scan<FK_BLOCK>(lq | (lu << 16), &tot2);
        uint32_t rq = e2 & 0xffffu, ru = e2 >> 16;
#pragma unroll
        for (int j = 0; j < SPT; j++) {
          const uint32_t c = hcnt[threadIdx.x * SPT + j];
          hpre[threadIdx.x * SPT + j] = rq | (ru << 16);
          rq += c;
          ru += (c + 31u) >> 5;
        }
        if (threadIdx.x == 0) { // the round trip of the reservation hides behind the query scatter below (result used after it)
          my_tu = tot2 >> 16;
          my_gu = my_tu ? atomicAdd(&ctl->unit_reserve, my_tu) : 0u;
        }
      }
      __syncthreads();
      FK_MARK(tl, 6); // slots scanned
      // The tile's queries in cell order: its range of the cell-ordered array IS its range of the compacted clouds (same
      // count, unique by the look-back), so only the unit list needs a reservation. One unit record per 32 queries of a slot.
      const uint32_t gq = s_prefix;
      uint32_t pos = gq;
      for (int w = 0; w < wid; w++)
        pos += wsum[w];
#pragma unroll
      for (int r = 0; r < FK_ROWS; r++) {
        if ((bal[r] >> lane) & 1u) {
          const uint32_t o = pos + __popc(bal[r] & lt);
          const uint32_t s = sr[r] & 0xffffu, rank = sr[r] >> 16;
          const float4 qm = tf_apply(T, pt[r]);
          q_sorted[gq + (hpre[s] & 0xffffu) + rank] = make_float4(qm.x, qm.y, qm.z, __int_as_float((int)o)); // tagged with its scan index
        }
        pos += __popc(bal[r]);
      }
      if (threadIdx.x == 0)
        s_gu = my_gu, s_tu = my_tu;
      __syncthreads();
      const uint32_t gu = s_gu, tu = s_tu;
      for (int s = threadIdx.x; s < FK_SLOTS; s += FK_BLOCK) {
        const uint32_t c = hcnt[s];
        if (c) {
          const uint32_t qb = gq + (hpre[s] & 0xffffu), ub = gu + (hpre[s] >> 16);
          int cx, cy, cz;
          cell_unkey(hkey[s], &cx, &cy, &cz);
          for (uint32_t j = 0; j * 32u < c; j++) {
            KnnUnit u;
            u.qbase = (int)(qb + 32u * j);
            u.cnt = (int)min(32u, c - 32u * j);
            u.cx = cx, u.cy = cy, u.cz = cz;
            units[ub + j] = u;
          }
        }
      }
      // publish: only the query and unit stores are outstanding at this fence (the clouds are written after it)
      __threadfence();
      __syncthreads();
      FK_MARK(tl, 7); // queries and unit records visible
      for (uint32_t j = threadIdx.x; j < tu; j += FK_BLOCK)
        units_tag[gu + j] = epoch;
      // the thread that reserved the units (thread 0) also counts the tile as done: its fence orders the reservation before
      // the count, which is all a leaving consumer relies on (unit_reserve final once tiles_done == ntiles)
      if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(&ctl->tiles_done, 1u);
      }
      // the clouds in scan order: pointcloud, pointcloud_original, pointcloud_in_map (nobody in this kernel reads them)
      pos = gq;
      for (int w = 0; w < wid; w++)
        pos += wsum[w];
#pragma unroll
      for (int r = 0; r < FK_ROWS; r++) {
        if ((bal[r] >> lane) & 1u) {
          const uint32_t o = pos + __popc(bal[r] & lt);
          const float4 q = make_float4(pt[r].x, pt[r].y, pt[r].z, t_rel);
          cloud[o] = q;
          orig[o] = q;
          in_map[o] = tf_apply(T, q);
        }
        pos += __popc(bal[r]);
      }
      if (next >= ntiles)
        break;
      tile = next;
      {
        const int nbase = (int)tile * RF_TILE + wid * (32 * FK_ROWS);
        if (VEC4) {
          mbar_wait(&s_bar, phase);
          phase ^= 1u;
        }
#pragma unroll
        for (int r = 0; r < FK_ROWS; r++) {
          const int i = nbase + r * 32 + lane;
          if (VEC4)
            pt[r] = i < n ? stage[wid * (32 * FK_ROWS) + r * 32 + lane] : make_float4(0.f, 0.f, 0.f, 0.f);
          else
            pt[r] = i < n ? load_raw_cg<VEC4>(raw, stride, i) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      __syncthreads(); // this tile's shared state is dead; the staged rays are in registers
    }
  }
  FK_MARK(tl, 2);

  // ------------------------------------------------------------------------------------------ consumer
  const int R = max(1, (int)ceil(sqrt((double)max_d2) * (1.0 + 1e-5) * g.inv_cell));
#if ML_KNN_STAGED
  // The candidate buffers of the 16 warps (6 KB each) take over the producer's dynamic shared memory: the hash table and the
  // stage are dead after a producer's last publication barrier, and a consumer-only CTA never used them.
  float4 *kbuf = reinterpret_cast<float4 *>(fk_smem) + (size_t)wid * (2 * KW_CH);
  uint32_t kpar = 0;
  if (lane == 0) {
    mbar_init(&s_kbar[wid][0], 1);
    mbar_init(&s_kbar[wid][1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
#endif
#if ML_KNN_PROF
  unsigned long long pw_wait = 0, pw_walk = 0, pw_emit = 0, pw_units = 0, pw_first = 0, pw_last = 0, pw_t = gtimer(), pw_rec = 0;
  unsigned long long pw_in[3] = {0, 0, 0};
#endif
  for (;;) {
    // A ticket is taken only when an unclaimed unit has been reserved (two relaxed loads, backing off up to 2 us while the
    // producers are still at work): thousands of warps spinning on tags and tickets cost the producers their L2 round trips
    // and their issue slots (a tile took 27 us instead of 8 us with every idle warp polling at 64 ns).
    uint32_t u = 0;
    bool have = false;
    if (lane == 0) {
      unsigned ns = 128;
      for (;;) {
        const uint32_t res = ld_relaxed_u32(&ctl->unit_reserve), nxt = ld_relaxed_u32(&ctl->unit_next);
        if (nxt < res) {
          u = atomicAdd(&ctl->unit_next, 1u);
          // wait for this unit to be published (it has been reserved, or will be within a tile's time); leave when every tile
          // is done and the ticket lies past the final count
          for (;;) {
            if (ld_acquire_u32(&units_tag[u]) == epoch) {
              have = true;
              break;
            }
            // every tile done: the reservation count is final; a ticket below it WILL be published, one past it never
            if (ld_acquire_u32(&ctl->tiles_done) == ntiles && u >= ld_acquire_u32(&ctl->unit_reserve))
              break;
            __nanosleep(256);
          }
          break;
        }
        if (ld_acquire_u32(&ctl->tiles_done) == ntiles && ld_relaxed_u32(&ctl->unit_next) >= ld_acquire_u32(&ctl->unit_reserve))
          break; // everything has been published and claimed
        __nanosleep(ns);
        ns = min(ns * 2u, 2048u);
      }
    }
    u = __shfl_sync(0xffffffffu, u, 0);
    have = __shfl_sync(0xffffffffu, (int)have, 0) != 0;
    __syncwarp();
    if (!have)
      break;
#if ML_KNN_PROF
    {
      const unsigned long long t = gtimer();
      pw_wait += t - pw_t, pw_t = t;
      if (!pw_first)
        pw_first = t;
    }
#endif
    // the unit record and its queries were written by another SM during this kernel: read them from L2 (an L1 line fetched for
    // an earlier unit may hold the bytes of this one from before they were written)
    KnnUnit un;
    {
      const int *up = reinterpret_cast<const int *>(units + u);
      un.qbase = __ldcg(up), un.cnt = __ldcg(up + 1), un.cx = __ldcg(up + 2), un.cy = __ldcg(up + 3), un.cz = __ldcg(up + 4);
    }
    const bool active = lane < un.cnt;
    const float4 q = __ldcg(&q_sorted[un.qbase + (active ? lane : 0)]);
    TopK<K> best;
#if ML_KNN_PROF
    {
      const unsigned long long t = gtimer() + (__float_as_int(q.x) == 0x7fffffff ? 1 : 0); // after the query has landed
      pw_rec += t - pw_t, pw_t = t;
    }
    knn_walk_staged<K>(un, q, active, g, R, cell_start, cell_pts, best, kbuf, s_kbar[wid], kpar, pw_in);
#elif ML_KNN_STAGED
    knn_walk_staged<K>(un, q, active, g, R, cell_start, cell_pts, best, kbuf, s_kbar[wid], kpar);
#else
    knn_walk<K>(un, q, active, g, R, cell_start, cell_pts, best);
#endif
#if ML_KNN_PROF
    {
      const unsigned long long t = gtimer();
      pw_walk += t - pw_t, pw_t = t;
    }
#endif
    if (active)
      knn_emit<K>(best, __float_as_int(q.w), max_d2, knn_idx, knn_d2, knn_ok, map_pts, nbr);
#if ML_KNN_PROF
    {
      __syncwarp();
      const unsigned long long t = gtimer();
      pw_emit += t - pw_t, pw_t = t, pw_last = t, pw_units++;
    }
#endif
  }
#if ML_KNN_PROF
  if (tl && lane == 0) {
    unsigned long long *o = tl + ((size_t)blockIdx.x * (FK_BLOCK / 32) + wid) * 8;
    o[0] = pw_wait, o[1] = pw_walk, o[2] = pw_emit, o[3] = pw_units | (pw_rec << 16), o[4] = pw_in[0] | (pw_in[1] << 20) | (pw_in[2] << 40), o[5] = pw_first, o[6] = pw_last, o[7] = (unsigned long long)((int)blockIdx.x < producers);
  }
#endif
  FK_MARK(tl, 3);
  // the last CTA to leave resets the control block for the next scan (every ticket has been taken by then)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&ctl->ctas_out, 1u) == gridDim.x - 1u) {
      ctl->tile_next = 0u, ctl->tiles_done = 0u, ctl->unit_reserve = 0u, ctl->unit_next = 0u;
      __threadfence();
      ctl->ctas_out = 0u;
    }
  }
}

} // namespace ml
