// Block/grid primitives used by the map and scan kernels: exclusive scan with a device-side element count,
// stable LSD radix sort of (key, value) pairs. Hand-written for sm_100a (no CUB/Thrust on the product path).
#pragma once
#include "common.cuh"

namespace ml {

constexpr int SCAN_BLOCK = 256;
constexpr int SCAN_ITEMS = 8, SCAN_ITEMS_LONG = 32; // per thread: short inputs (points of a scan / of the map), long inputs (grid cells)
constexpr int SCAN_TILE = SCAN_BLOCK * SCAN_ITEMS, SCAN_TILE_LONG = SCAN_BLOCK * SCAN_ITEMS_LONG;

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o)
      v += t;
  }
  return v;
}

// exclusive scan of one value per thread across the block; returns the exclusive prefix, *total = block sum.
template <int BLOCK> __device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *total) {
  __shared__ uint32_t wsum[BLOCK / 32];
  __shared__ uint32_t btot;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t inc = warp_incl_scan(v);
  if (lane == 31)
    wsum[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    uint32_t s = lane < BLOCK / 32 ? wsum[lane] : 0;
    uint32_t si = warp_incl_scan(s);
    if (lane < BLOCK / 32)
      wsum[lane] = si - s;
    if (lane == 31)
      btot = si;
  }
  __syncthreads();
  uint32_t r = wsum[wid] + inc - v;
  if (total)
    *total = btot;
  __syncthreads();
  return r;
}

// ------------------------------------------------------------------------------------------------
// Decoupled look-back (single-pass prefix over tiles). A tile publishes (epoch, AGG, its count), walks backwards over its
// predecessors 32 at a time adding counts until it meets one that already knows its inclusive prefix, then publishes
// (epoch, PREFIX, inclusive prefix). Tiles are handed out by an atomic ticket, so every tile a CTA waits for belongs to a CTA
// that is already running: no deadlock whatever the grid size. States carry the launch's epoch: no reset pass between
// launches (a state of another epoch reads as "not published yet"). Only CTAs that own a tile take a ticket. Integer sums: the result does not depend on timing.
// ------------------------------------------------------------------------------------------------
constexpr unsigned long long LB_AGG = 1ull, LB_PREFIX = 2ull;
__device__ __forceinline__ unsigned long long lb_pack(uint32_t epoch, unsigned long long flag, uint32_t v) {
  return ((unsigned long long)epoch << 34) | (flag << 32) | (unsigned long long)v;
}
// tile ticket for the `active` CTAs of a launch (exactly that many call it); the counter is back at 0 when all are handed out
__device__ __forceinline__ uint32_t lb_ticket(unsigned int *tile_ctr, uint32_t active) {
  const uint32_t t = atomicAdd(tile_ctr, 1u);
  if (t == active - 1u)
    atomicExch(tile_ctr, 0u);
  return t;
}
// warp-collective (all 32 lanes of one warp of the CTA that owns `tile`); returns the exclusive prefix in every lane
__device__ __forceinline__ uint32_t lb_exclusive_prefix(unsigned long long *tile_state, uint32_t tile, uint32_t epoch, uint32_t tot) {
  const int lane = threadIdx.x & 31;
  // a state is ONE 64-bit word (epoch, flag, value) and readers consume nothing else the publisher wrote: no fence is
  // needed for correctness; the atomic is there because it is performed at L2 at once (a plain store is visible later)
  if (lane == 0)
    atomicExch(&tile_state[tile], lb_pack(epoch, tile == 0 ? LB_PREFIX : LB_AGG, tot));
  uint32_t excl = 0;
  if (tile > 0) {
    int base = (int)tile - 1;
    for (;;) {
      const int b = base - lane;
      unsigned long long sv = lb_pack(epoch, LB_PREFIX, 0u); // tiles before the first: prefix 0
      if (b >= 0) {
        do {
          asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(sv) : "l"(tile_state + b) : "memory");
        } while ((uint32_t)(sv >> 34) != epoch);
      }
      const bool is_prefix = ((sv >> 32) & 3ull) == LB_PREFIX;
      const uint32_t val = (uint32_t)sv;
      const unsigned pm = __ballot_sync(0xffffffffu, is_prefix);
      // lanes up to (and including) the nearest tile that already knows its prefix contribute
      const int first = pm ? __ffs(pm) - 1 : 31;
      uint32_t part = lane <= first ? val : 0u;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
        part += __shfl_xor_sync(0xffffffffu, part, o);
      excl += part;
      if (pm)
        break;
      base -= 32;
    }
    if (lane == 0)
      atomicExch(&tile_state[tile], lb_pack(epoch, LB_PREFIX, excl + tot));
  }
  return excl;
}

// ------------------------------------------------------------------------------------------------
// Exclusive scan of uint32 with a device-side element count: out[i] = sum(in[0..i)), out[n] = total. n is read from device
// memory so a launch can be sized by a host-side upper bound; CTAs whose tile starts past n leave at once.
// A CTA tile is SCAN_BLOCK*ITEMS elements, warp-striped: warp w owns 32*ITEMS consecutive elements, a row is 128 of them,
// a lane holds 4 consecutive ones (one 16-byte access; a warp instruction covers 512 contiguous bytes).
// ZERO_IN: non-zero inputs are zeroed once read (the kNN grid's cell counters: all-zero again after the build, without a
// clearing pass over the mostly empty dense grid).
// Driver: k_scan_lb (one pass, decoupled look-back). scripts/bench_scan.cu holds the two-pass alternative (reduce + apply on the
// same tile code) it was measured against: 37 us vs 30 us on the 13.5 M cell counters.
// ------------------------------------------------------------------------------------------------
template <int ITEMS, bool ZERO_IN> struct ScanTile {
  static constexpr int ROWS = ITEMS / 4, TILE = SCAN_BLOCK * ITEMS;
  uint4 q[ROWS];
  uint32_t ex[ROWS]; // exclusive prefix, inside the CTA tile, of the lane's 4 items of row r

  __device__ __forceinline__ void load(uint32_t *in, long long base, int n) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long long wbase = base + (long long)wid * (32 * ITEMS);
#pragma unroll
    for (int r = 0; r < ROWS; r++) {
      const long long i = wbase + r * 128 + lane * 4;
      if (i + 4 <= (long long)n) {
        q[r] = *reinterpret_cast<const uint4 *>(in + i);
        if (ZERO_IN && (q[r].x | q[r].y | q[r].z | q[r].w))
          *reinterpret_cast<uint4 *>(in + i) = make_uint4(0u, 0u, 0u, 0u);
      } else {
        uint32_t v[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
          v[k] = i + k < (long long)n ? in[i + k] : 0u;
          if (ZERO_IN && v[k])
            in[i + k] = 0u;
        }
        q[r] = make_uint4(v[0], v[1], v[2], v[3]);
      }
    }
  }
  // sum of the CTA tile (every thread)
  __device__ __forceinline__ uint32_t reduce() {
    __shared__ uint32_t wsum[SCAN_BLOCK / 32];
    uint32_t s = 0;
#pragma unroll
    for (int r = 0; r < ROWS; r++)
      s += q[r].x + q[r].y + q[r].z + q[r].w;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0)
      wsum[threadIdx.x >> 5] = s;
    __syncthreads();
    uint32_t tot = 0;
#pragma unroll
    for (int w = 0; w < SCAN_BLOCK / 32; w++)
      tot += wsum[w];
    return tot;
  }
  // fills ex[]; returns the sum of the CTA tile (every thread)
  __device__ __forceinline__ uint32_t scan() {
    __shared__ uint32_t wsum[SCAN_BLOCK / 32];
    const int wid = threadIdx.x >> 5;
    uint32_t off = 0;
#pragma unroll
    for (int r = 0; r < ROWS; r++) {
      const uint32_t rs = q[r].x + q[r].y + q[r].z + q[r].w;
      const uint32_t inc = warp_incl_scan(rs);
      ex[r] = off + inc - rs;
      off += __shfl_sync(0xffffffffu, inc, 31);
    }
    if ((threadIdx.x & 31) == 0)
      wsum[wid] = off;
    __syncthreads();
    uint32_t wpre = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < SCAN_BLOCK / 32; w++) {
      wpre += w < wid ? wsum[w] : 0u;
      tot += wsum[w];
    }
#pragma unroll
    for (int r = 0; r < ROWS; r++)
      ex[r] += wpre;
    return tot;
  }
  __device__ __forceinline__ void store(uint32_t *out, long long base, int n, uint32_t prefix) const {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long long wbase = base + (long long)wid * (32 * ITEMS);
#pragma unroll
    for (int r = 0; r < ROWS; r++) {
      const long long i = wbase + r * 128 + lane * 4;
      uint4 a;
      a.x = prefix + ex[r], a.y = a.x + q[r].x, a.z = a.y + q[r].y, a.w = a.z + q[r].z;
      if (i + 4 <= (long long)n) {
        *reinterpret_cast<uint4 *>(out + i) = a;
      } else {
        const uint32_t v[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int k = 0; k < 4; k++)
          if (i + k < (long long)n)
            out[i + k] = v[k];
      }
    }
  }
};

template <int ITEMS, bool ZERO_IN>
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_lb(uint32_t *in, const int *__restrict__ d_n, uint32_t *__restrict__ out,
                                                       int *__restrict__ d_total, unsigned long long *tile_state,
                                                       unsigned int *tile_ctr, uint32_t epoch) {
  constexpr int TILE = SCAN_BLOCK * ITEMS;
  __shared__ uint32_t s_tile, s_prefix;
  const int n = *d_n;
  const uint32_t active = n > 0 ? (uint32_t)((n + TILE - 1) / TILE) : 1u;
  if (blockIdx.x >= active)
    return;
  if (threadIdx.x == 0)
    s_tile = lb_ticket(tile_ctr, active);
  __syncthreads();
  const uint32_t tile = s_tile;
  const long long base = (long long)tile * TILE;
  ScanTile<ITEMS, ZERO_IN> t;
  t.load(in, base, n);
  const uint32_t tot = t.scan();
  if (threadIdx.x < 32) {
    const uint32_t excl = lb_exclusive_prefix(tile_state, tile, epoch, tot);
    if (threadIdx.x == 0)
      s_prefix = excl;
  }
  __syncthreads();
  t.store(out, base, n, s_prefix);
  if (base + TILE >= (long long)n && threadIdx.x == 0) {
    out[n] = s_prefix + tot;
    if (d_total)
      *d_total = (int)(s_prefix + tot);
  }
}

// ------------------------------------------------------------------------------------------------
// Stable LSD radix sort, 8 bits per pass, of (key uint32, value uint32) pairs; n on the device.
// One histogram for all passes, built by the kernel that produces the keys (radix_hist4_add / radix_hist4_publish, used by
// k_vox_index_hist: the CTA that finishes last turns the four digit histograms into digit bases), then ONE kernel per pass (k_radix_pass): stable in-tile ranking with warp match_any
// (warps own consecutive chunks), and the tile's offset inside every digit's output range by a decoupled look-back per digit
// (thread d follows digit d through the preceding tiles). 5 launches for a 32-bit sort instead of 12.
// ------------------------------------------------------------------------------------------------
constexpr int RS_BLOCK = 256;
constexpr int RS_ITEMS = 4; // keys per thread (16 measured no faster on the 1 M-point map: the in-tile ranking dominates, not the look-back)
constexpr int RS_TILE = RS_BLOCK * RS_ITEMS; // 1024: a 131 072-point scan sorts on 128 CTAs
constexpr int RS_WARPS = RS_BLOCK / 32;
constexpr int RS_PASSES = 4;

// Digit histograms of all passes at once. h[p][d] (shared) = this CTA's counts. ghist[p][d] (global, all-zero on entry and on
// exit) accumulates them; the CTA that finishes last writes dbase[p][d] = number of keys whose p-th byte is smaller than d.
// `done` is 0 on entry and on exit. Called by every thread of the `active` CTAs after their shared histogram is complete.
__device__ __forceinline__ void radix_hist4_publish(uint32_t (*h)[256], uint32_t *ghist, uint32_t *__restrict__ dbase,
                                                    unsigned int *done, uint32_t active) {
  __shared__ uint32_t s_last;
  __syncthreads();
#pragma unroll
  for (int p = 0; p < RS_PASSES; p++)
    if (h[p][threadIdx.x])
      atomicAdd(&ghist[p * 256 + threadIdx.x], h[p][threadIdx.x]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0)
    s_last = atomicAdd(done, 1u) == active - 1u;
  __syncthreads();
  if (!s_last)
    return;
  __threadfence();
#pragma unroll
  for (int p = 0; p < RS_PASSES; p++) {
    const uint32_t v = *(volatile uint32_t *)&ghist[p * 256 + threadIdx.x];
    ghist[p * 256 + threadIdx.x] = 0u;
    dbase[p * 256 + threadIdx.x] = block_excl_scan<RS_BLOCK>(v, nullptr);
  }
  if (threadIdx.x == 0)
    *done = 0u;
}
// plain shared-memory atomics: warp-aggregating them with match_any (the high bytes of voxel keys take few values) was
// measured slower (11.0 vs 9.6 us for 131 072 keys)
__device__ __forceinline__ void radix_hist4_add(uint32_t (*h)[256], uint32_t key) {
#pragma unroll
  for (int p = 0; p < RS_PASSES; p++)
    atomicAdd(&h[p][(key >> (8 * p)) & 255u], 1u);
}

template <int ITEMS>
__global__ void __launch_bounds__(RS_BLOCK) k_radix_pass(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals,
                                                        const int *__restrict__ d_n, int pass, const uint32_t *__restrict__ dbase,
                                                        unsigned long long *digit_state, unsigned int *tile_ctr, uint32_t epoch,
                                                        uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out) {
  __shared__ uint32_t wcnt[RS_WARPS][256]; // per-warp digit counts -> per-warp bases
  __shared__ uint32_t s_tile;
  constexpr int TILE = RS_BLOCK * ITEMS;
  const int n = *d_n;
  const uint32_t active = n > 0 ? (uint32_t)((n + TILE - 1) / TILE) : 1u;
  if (blockIdx.x >= active)
    return;
  const int shift = 8 * pass;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0)
    s_tile = lb_ticket(tile_ctr, active);
  for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_BLOCK)
    (&wcnt[0][0])[i] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  // warp w owns the contiguous chunk [base + w*CH, base + (w+1)*CH), processed 32 at a time in order
  constexpr int CH = TILE / RS_WARPS;
  const int wbase = (int)tile * TILE + wid * CH;
  uint32_t myk[CH / 32], myv[CH / 32], myrank[CH / 32];
#pragma unroll
  for (int r = 0; r < CH / 32; r++) {
    const int i = wbase + r * 32 + lane;
    myk[r] = i < n ? keys[i] : 0xffffffffu;
    myv[r] = i < n ? vals[i] : 0u;
  }
#pragma unroll
  for (int r = 0; r < CH / 32; r++) {
    const int i = wbase + r * 32 + lane;
    const bool ok = i < n;
    const uint32_t d = (myk[r] >> shift) & 255u;
    const unsigned act = __ballot_sync(0xffffffffu, ok);
    uint32_t rank = 0;
    if (ok) {
      const unsigned peers = __match_any_sync(act, d);
      rank = wcnt[wid][d] + __popc(peers & ((1u << lane) - 1u));
      __syncwarp(act);
      if ((peers & ((1u << lane) - 1u)) == 0)
        wcnt[wid][d] += __popc(peers);
    }
    __syncwarp();
    myrank[r] = rank;
  }
  __syncthreads();
  {
    // thread d: this tile's count of digit d, its offset inside digit d's output range by look-back over the preceding
    // tiles (state word = {epoch, flag, value}, published with an L2 atomic: see lb_exclusive_prefix), per-warp bases
    const int d = threadIdx.x; // 256 threads == 256 digits
    uint32_t cnt = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++)
      cnt += wcnt[w][d];
    atomicExch(&digit_state[(size_t)tile * 256 + d], lb_pack(epoch, tile == 0 ? LB_PREFIX : LB_AGG, cnt));
    uint32_t excl = 0;
    // LB_WALK predecessors per round trip (their loads are in flight together); consumed in order, a state that is not
    // published yet is re-read until it is
    constexpr int LB_WALK = 8;
    bool done_walk = false;
    for (int b = (int)tile - 1; b >= 0 && !done_walk; b -= LB_WALK) {
      unsigned long long sv[LB_WALK];
#pragma unroll
      for (int q = 0; q < LB_WALK; q++)
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(sv[q]) : "l"(digit_state + (size_t)max(b - q, 0) * 256 + d) : "memory");
#pragma unroll
      for (int q = 0; q < LB_WALK; q++) {
        if (done_walk || b - q < 0)
          continue;
        unsigned long long v = sv[q];
        while ((uint32_t)(v >> 34) != epoch)
          asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(digit_state + (size_t)(b - q) * 256 + d) : "memory");
        excl += (uint32_t)v;
        done_walk = ((v >> 32) & 3ull) == LB_PREFIX;
      }
    }
    if (tile > 0)
      atomicExch(&digit_state[(size_t)tile * 256 + d], lb_pack(epoch, LB_PREFIX, excl + cnt));
    uint32_t run = dbase[pass * 256 + d] + excl;
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) {
      const uint32_t c = wcnt[w][d];
      wcnt[w][d] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < CH / 32; r++) {
    const int i = wbase + r * 32 + lane;
    if (i < n) {
      const uint32_t d = (myk[r] >> shift) & 255u;
      const uint32_t pos = wcnt[wid][d] + myrank[r];
      keys_out[pos] = myk[r];
      vals_out[pos] = myv[r];
    }
  }
}

} // namespace ml
