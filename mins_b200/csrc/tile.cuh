// The whole un-down-sampled scan update as ONE kernel in which a thread owns a ray from the raw point to its row of the linear
// system, and a CTA owns a tile of 448 consecutive rays down to the tile's share of the reduced system. Replaces k_front_knn +
// k_rows_reduce (2 launches, a shared-memory hash grouping per tile, a global unit queue with tickets, tags and fences, and the
// q_sorted / units / nbr round trips through L2) on the one-call path.
//
// What the measurements said about the producer / consumer kernel (profiles/r2_ncu.md, scripts/knn_warp_profile.py, G-64):
//   * a warp-per-cell unit of the search holds 15 queries and 26 candidates on average: its cost is not arithmetic but a chain
//     of dependent round trips (ticket -> tag -> unit record -> query -> cell_start -> candidates -> neighbour coordinates),
//     8.6 us per unit, two units per warp; the search phase lasted 25 us for 13 us worth of issue slots, and more than half of
//     the 14.8 M warp instructions were queueing, grouping and set-up, at 15 of 32 lanes active;
//   * the 64 tiles of the front end finished together after 14 us, during which 232 of the 296 CTAs polled;
//   * the rows kernel started 4 us after the last search CTA had left, spent 2 us on its own prologue, and its first
//     instruction needed the neighbour coordinates back from L2.
// Here nothing is grouped and nothing is queued. Tile t belongs to CTA t mod grid (static: the reduction order is a function
// of the scan size only, results are bit-reproducible); 2 CTAs per SM, so a 131 072-ray scan is one wave. Per tile:
//   1 the tile's rays by one TMA bulk copy (from pinned host memory the copies of all resident CTAs are in flight together and
//     a CTA goes to work when its own tile has landed), range gate (LidarTypes.cpp:36-55), ballots; warp 0 starts the
//     look-back for the tile's offset in the compacted clouds (needed for the OUTPUT clouds only);
//   2 transform (LidarHelper.cpp:192) and the exact kNN of the thread's own query: knn_walk_lane (update.cuh) - the cube walk
//     with the arithmetic, completeness rule and (d2, index) order of the grouped walk, hence the same neighbours; lanes of a
//     warp whose queries share a cell coalesce, the rest is served by L1 (a tile's cubes overlap heavily);
//   3 neighbour coordinates gathered from the map, plane fit, residual, folded Jacobian, chi2 gate (point_row, rows.cuh); the
//     warp's 32 rows reduced through its shared-memory slab (phase 2 of k_rows_reduce), the CTA's 14 warp partials added in
//     warp order into the CTA's running system;
//   4 clouds (pointcloud, pointcloud_original, pointcloud_in_map) and, on request (MINS_TILE_NBR=1), the neighbours' coordinates in
//     scan order; by default ml_point_rows regenerates them with the multi-call search before it re-runs the rows export.
// At the end a CTA writes its reduced system into its RRPart in pinned host memory; the host, which armed the parts with a
// sentinel pattern, adds them in CTA order as their words turn up (rows_host_tail): no stream synchronisation, no flag, no fence.
#pragma once
#include "fused.cuh"
#include "rows.cuh"
#include "update.cuh"

#ifndef ML_TILE_PREFETCH
#define ML_TILE_PREFETCH 0 // see knn_walk_lane: measured slower
#endif

namespace ml {

constexpr int ST_BLOCK = 448;           // one ray per thread; 2 CTAs per SM at <= 72 registers
constexpr int ST_WARPS = ST_BLOCK / 32; // 14
constexpr int ST_TILE = ST_BLOCK;

template <int NB> struct STLayout {
  static constexpr int NG = 3 * NB, ROW = NG + 2, ROLES = NB * (NB + 1) / 2;
  static constexpr int NSYM = NG * (NG + 1) / 2, NE = NSYM + NG, NRED = NE + 4;
  // per warp: g rows (odd stride) | W | V | the warp's partial (ROLES x RR_RACC) | 4 statistics
  static constexpr int OFF_W = 32 * ROW, OFF_V = OFF_W + 32, OFF_PART = OFF_V + 32, OFF_STAT = OFF_PART + ROLES * RR_RACC;
  static constexpr int SLAB = OFF_STAT + 4; // doubles
  static constexpr size_t BYTES = sizeof(double) * SLAB * ST_WARPS;
  static_assert(BYTES >= sizeof(float4) * ST_TILE, "the raw tile is staged in the slabs");
};

template <int K, int NB, bool VEC4>
__global__ void __launch_bounds__(ST_BLOCK, 2)
    k_scan_tile(const __grid_constant__ RowConsts rc, const float *__restrict__ raw, int stride, int n, float t_rel, double rmin,
                double rmax, ScanState *__restrict__ st, float4 *__restrict__ cloud, float4 *__restrict__ orig, Tf32 T,
                float4 *__restrict__ in_map, unsigned long long *tile_state, uint32_t epoch, const GridDesc *__restrict__ gd,
                const uint32_t *__restrict__ cell_start, const float4 *__restrict__ cell_pts, const float4 *__restrict__ map_pts,
                float max_d2, float4 *__restrict__ nbr, RRPart *__restrict__ parts /* pinned host, armed */,
                unsigned long long *__restrict__ tl) {
  using LY = STLayout<NB>;
  constexpr int NG = LY::NG, ROW = LY::ROW, ROLES = LY::ROLES, NSYM = LY::NSYM, NE = LY::NE, NRED = LY::NRED;
  extern __shared__ __align__(16) unsigned char st_smem[];
  double *slabs = reinterpret_cast<double *>(st_smem);
  float4 *stage = reinterpret_cast<float4 *>(st_smem);
  __shared__ __align__(8) unsigned long long s_bar; // the tile's bulk copy
  __shared__ uint32_t wsum[ST_WARPS];
  __shared__ uint32_t s_prefix;
  __shared__ double s_cta[ML_NRED]; // the CTA's running reduced system
  __shared__ int s_cnt[5];
  __shared__ int s_nscan;
  tl_mark(tl, 0);
  pdl_enter();
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t lt = (1u << lane) - 1u;
  const uint32_t ntiles = (uint32_t)((n + ST_TILE - 1) / ST_TILE);
  if (VEC4 && tid == 0) {
    mbar_init(&s_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // the first tile's rays start their journey before anything else is set up
    const uint32_t bytes = (uint32_t)min(ST_TILE, n - (int)blockIdx.x * ST_TILE) * (uint32_t)sizeof(float4);
    mbar_expect_tx(&s_bar, bytes);
    bulk_g2s(stage, reinterpret_cast<const float4 *>(raw) + (size_t)blockIdx.x * ST_TILE, bytes, &s_bar);
  }
  const GridDesc g = *gd;
  const int R = max(1, (int)ceil(sqrt((double)max_d2) * (1.0 + 1e-5) * g.inv_cell));
  if (tid == 0)
    s_nscan = -1;
  if (tid < NRED)
    s_cta[tid] = 0.0;
  if (tid < 5)
    s_cnt[tid] = 0;
  __syncthreads();
  uint32_t tphase = 0;

  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int i = (int)tile * ST_TILE + tid;
    // ------------------------------------------------------------------ 1: the tile's rays
    if (VEC4 && tile != blockIdx.x && tid == 0) {
      const uint32_t bytes = (uint32_t)min(ST_TILE, n - (int)tile * ST_TILE) * (uint32_t)sizeof(float4);
      mbar_expect_tx(&s_bar, bytes);
      bulk_g2s(stage, reinterpret_cast<const float4 *>(raw) + (size_t)tile * ST_TILE, bytes, &s_bar);
    }
    float4 pt;
    if (VEC4) {
      // one thread watches the mbarrier, the others sleep in the hardware barrier: with the scan in pinned host memory a CTA
      // waits tens of microseconds for its tile while its neighbour on the SM is already at work and wants the issue slots
      if (tid == 0)
        mbar_wait(&s_bar, tphase);
      tphase ^= 1u;
      __syncthreads();
      pt = i < n ? stage[tid] : make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      pt = i < n ? load_raw_cg<false>(raw, stride, i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const bool kept = i < n && range_keep(pt, rmin, rmax);
    const uint32_t bal = __ballot_sync(0xffffffffu, kept);
    if (lane == 0)
      wsum[wid] = __popc(bal);
    __syncthreads(); // counts known; the staged rays are in registers (the slabs are free)
    tl_mark(tl, 1);
    // warp 0: the tile's offset in the compacted clouds (decoupled look-back); nobody needs it before the stores at the end
    if (wid == 0) {
      uint32_t tot = lane < ST_WARPS ? wsum[lane] : 0u;
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
          s_nscan = (int)(excl + tot);
        }
      }
    }
    // ------------------------------------------------------------------ 2: the query's exact k nearest map points
    const float4 qm = tf_apply(T, pt);
    TopK<K> best;
    best.init();
    if (kept) {
      const int lim = 1 << 19;
      const int cx = max(-lim, min(lim, cell_coord((double)qm.x, g.ox, g.inv_cell)));
      const int cy = max(-lim, min(lim, cell_coord((double)qm.y, g.oy, g.inv_cell)));
      const int cz = max(-lim, min(lim, cell_coord((double)qm.z, g.oz, g.inv_cell)));
      knn_walk_lane<K, ML_TILE_PREFETCH != 0>(qm, cx, cy, cz, g, R, cell_start, cell_pts, best);
    }
    // get_neighbors' verdict (LidarHelper.cpp:280-285): k neighbours, the farthest within map_ngbr_max_d (squared distance
    // against the unsquared option, as the reference has it)
    const bool has_nb = kept && best.i[K - 1] != 0x7fffffff && !((double)best.d[K - 1] > (double)max_d2);
    tl_mark(tl, 2);
    // ------------------------------------------------------------------ 3: the ray's row, the warp's and the CTA's sums
    double *sg = slabs + (size_t)wid * LY::SLAB; // [32][ROW]
    double *sw = sg + LY::OFF_W, *sv = sg + LY::OFF_V, *spart = sg + LY::OFF_PART, *sstat = sg + LY::OFF_STAT;
    float4 nb[K];
    {
      int status = 1;
      double gv[NG];
      RowOut ro;
      ro.W = ro.V = ro.chi2 = ro.rss = ro.raw_n = ro.res5 = ro.d = ro.sum_rp = 0.0;
      ro.n[0] = ro.n[1] = ro.n[2] = 0.0;
#pragma unroll
      for (int a = 0; a < NG; a++)
        gv[a] = 0.0;
#pragma unroll
      for (int j = 0; j < K; j++)
        nb[j] = has_nb ? __ldg(&map_pts[best.i[j]]) : make_float4(0.f, 0.f, 0.f, 0.f);
      if (has_nb)
        status = point_row<K, NB, false>(rc, make_float4(pt.x, pt.y, pt.z, t_rel), nb, gv, ro);
      const bool okp = kept && status == 0;
#pragma unroll
      for (int a = 0; a < NG; a++)
        sg[lane * ROW + a] = okp ? gv[a] : 0.0;
      sw[lane] = okp ? ro.W : 0.0;
      sv[lane] = okp ? ro.V : 0.0;
      const bool tested = kept && (status == 0 || status == 4) && (ro.chi2 == ro.chi2);
      double vchi = tested ? ro.chi2 : 0.0, vrss = tested ? ro.rss : 0.0;
      double vchi2 = vchi * vchi;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        vchi += __shfl_xor_sync(0xffffffffu, vchi, o);
        vchi2 += __shfl_xor_sync(0xffffffffu, vchi2, o);
        vrss += __shfl_xor_sync(0xffffffffu, vrss, o);
      }
      if (lane == 0)
        sstat[0] = vchi, sstat[1] = vchi2, sstat[2] = vrss, sstat[3] = 0.0;
      int mine = 0;
#pragma unroll
      for (int q = 0; q < 5; q++) {
        const int cq = __popc(__ballot_sync(0xffffffffu, kept && status == q));
        mine = lane == q ? cq : mine;
      }
      if (lane < 5 && mine)
        atomicAdd(&s_cnt[lane], mine);
    }
    __syncwarp();
    // phase 2 of k_rows_reduce: lane = (group of points, 3x3 block (bi, bj) of G); the groups are folded by shuffles in group
    // order and the warp's partial (ROLES x 12 numbers) is left in its slab
    {
      constexpr int GROUPS = 32 / ROLES;
      const int grp = lane / ROLES, role = lane - grp * ROLES;
      const bool p2_lane = lane < ROLES * GROUPS;
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
      double acc[RR_RACC];
#pragma unroll
      for (int q = 0; q < RR_RACC; q++)
        acc[q] = 0.0;
      if (p2_lane) {
        const double *gI = sg + 3 * bi, *gJ = sg + 3 * bj;
#pragma unroll 4
        for (int p = grp; p < 32; p += GROUPS) {
          const double w = sw[p];
          const double x0 = gI[p * ROW] * w, x1 = gI[p * ROW + 1] * w, x2 = gI[p * ROW + 2] * w;
          const double y0 = gJ[p * ROW], y1 = gJ[p * ROW + 1], y2 = gJ[p * ROW + 2];
          acc[0] += x0 * y0, acc[1] += x0 * y1, acc[2] += x0 * y2;
          acc[3] += x1 * y0, acc[4] += x1 * y1, acc[5] += x1 * y2;
          acc[6] += x2 * y0, acc[7] += x2 * y1, acc[8] += x2 * y2;
          if (bi == bj) {
            const double v = sv[p];
            acc[9] += y0 * v, acc[10] += y1 * v, acc[11] += y2 * v;
          }
        }
      }
#pragma unroll
      for (int q = 0; q < RR_RACC; q++) {
        double tot = acc[q];
#pragma unroll
        for (int gq = 1; gq < GROUPS; gq++)
          tot += __shfl_sync(0xffffffffu, acc[q], (lane + gq * ROLES) & 31);
        if (lane < ROLES)
          spart[lane * RR_RACC + q] = tot;
      }
    }
    __syncthreads();
    tl_mark(tl, 3);
    // the CTA's running system: entry e += the 14 warp partials in warp order
    if (tid < NE) {
      int erole, q;
      if (tid < NSYM) {
        int r = 0, rem = tid;
        while (rem >= NG - r) {
          rem -= NG - r;
          r++;
        }
        const int gi = r, gj = r + rem;
        erole = rr_role_of<NB>(gi / 3, gj / 3);
        q = (gi % 3) * 3 + (gj % 3);
      } else {
        const int a = tid - NSYM;
        erole = rr_role_of<NB>(a / 3, a / 3);
        q = 9 + a % 3;
      }
      double v = 0.0;
      for (int w = 0; w < ST_WARPS; w++)
        v += slabs[(size_t)w * LY::SLAB + LY::OFF_PART + erole * RR_RACC + q];
      s_cta[tid] += v;
    } else if (tid < NRED) {
      double v = 0.0;
      for (int w = 0; w < ST_WARPS; w++)
        v += slabs[(size_t)w * LY::SLAB + LY::OFF_STAT + (tid - NE)];
      s_cta[tid] += v;
    }
    // The CTA's last tile: the reduced system goes out BEFORE the clouds, as plain posted writes into pinned host memory. There
    // is no flag and no fence: the host armed every word of the part with a sentinel bit pattern before the launch and takes a
    // word as soon as it reads anything else (aligned 8- and 4-byte stores arrive whole). A system-scope fence in front of a
    // stamp cost 7 us per CTA here (10 us with one fence per writing thread).
    if (tile + gridDim.x >= ntiles) {
      RRPart *mine = parts + blockIdx.x;
      if (tid < NRED)
        mine->v[tid] = s_cta[tid];
      else if (tid >= 192 && tid < 197)
        mine->c[tid - 192] = s_cnt[tid - 192];
      else if (tid == 197)
        mine->n_scan = s_nscan;
      tl_mark(tl, 4);
    }
    // ------------------------------------------------------------------ 4: the clouds in scan order
    if (kept) {
      uint32_t o = s_prefix + (uint32_t)__popc(bal & lt);
      for (int w = 0; w < wid; w++)
        o += wsum[w];
      const float4 q4 = make_float4(pt.x, pt.y, pt.z, t_rel);
      cloud[o] = q4;
      orig[o] = q4;
      in_map[o] = tf_apply(T, q4);
      if (nbr) {
        if (has_nb) {
#pragma unroll
          for (int j = 0; j < K; j++)
            nbr[(size_t)o * K + j] = nb[j];
        } else {
          nbr[(size_t)o * K] = make_float4(__int_as_float(0x7fc00000), 0.f, 0.f, 0.f);
        }
      }
    }
    __syncthreads(); // the slabs, wsum and s_prefix are free for the next tile
  }
  tl_mark(tl, 5);
}

} // namespace ml
