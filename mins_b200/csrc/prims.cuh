// Block/grid primitives used by the map and scan kernels: exclusive scan with a device-side element count,
// stable LSD radix sort of (key, value) pairs. Hand-written for sm_100a (no CUB/Thrust on the product path).
#pragma once
#include "common.cuh"

namespace ml {

constexpr int SCAN_BLOCK = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_BLOCK * SCAN_ITEMS;

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

// pass 1: per-tile sums. n is read from device memory so the launch can be sized by capacity.
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_reduce(const uint32_t *__restrict__ in, const int *__restrict__ d_n,
                                                           uint32_t *__restrict__ tile_sums) {
  const int n = *d_n;
  const int base = blockIdx.x * SCAN_TILE;
  if (base >= n)
    return;
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    int i = base + k * SCAN_BLOCK + threadIdx.x;
    if (i < n)
      s += in[i];
  }
  uint32_t tot;
  block_excl_scan<SCAN_BLOCK>(s, &tot);
  if (threadIdx.x == 0)
    tile_sums[blockIdx.x] = tot;
}

// pass 2: out[i] = sum(in[0..i)); out[n] = total. Each tile sums the preceding tile sums itself (fixed order,
// no inter-block dependency, deterministic).
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_apply(const uint32_t *__restrict__ in, const int *__restrict__ d_n,
                                                          const uint32_t *__restrict__ tile_sums, uint32_t *__restrict__ out,
                                                          int *__restrict__ d_total) {
  const int n = *d_n;
  const int base = blockIdx.x * SCAN_TILE;
  if (base >= n) {
    if (n == 0 && blockIdx.x == 0 && threadIdx.x == 0) {
      out[0] = 0;
      if (d_total)
        *d_total = 0;
    }
    return;
  }
  uint32_t p = 0;
  for (int b = threadIdx.x; b < blockIdx.x; b += SCAN_BLOCK)
    p += tile_sums[b];
  uint32_t prefix;
  block_excl_scan<SCAN_BLOCK>(p, &prefix);
  // blocked arrangement: thread t owns items [t*ITEMS, t*ITEMS+ITEMS)
  uint32_t v[SCAN_ITEMS], s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    int i = base + threadIdx.x * SCAN_ITEMS + k;
    v[k] = i < n ? in[i] : 0;
    s += v[k];
  }
  uint32_t tot;
  uint32_t ex = block_excl_scan<SCAN_BLOCK>(s, &tot) + prefix;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    int i = base + threadIdx.x * SCAN_ITEMS + k;
    if (i < n)
      out[i] = ex;
    ex += v[k];
  }
  if (base + SCAN_TILE >= n && threadIdx.x == 0) {
    out[n] = prefix + tot;
    if (d_total)
      *d_total = (int)(prefix + tot);
  }
}

// ------------------------------------------------------------------------------------------------
// Stable LSD radix sort, 8 bits per pass, of (key uint32, value uint32) pairs; n on the device.
// Per pass: k_radix_hist (tile histograms, digit-major table) -> k_radix_scan (one block scans the
// table) -> k_radix_scatter (stable in-tile ranking with warp match_any, warps own consecutive chunks).
// ------------------------------------------------------------------------------------------------
constexpr int RS_BLOCK = 256;
constexpr int RS_ITEMS = 4;
constexpr int RS_TILE = RS_BLOCK * RS_ITEMS; // 1024: a 131 072-point scan sorts on 128 CTAs
constexpr int RS_WARPS = RS_BLOCK / 32;

__global__ void __launch_bounds__(RS_BLOCK) k_radix_hist(const uint32_t *__restrict__ keys, const int *__restrict__ d_n, int shift,
                                                        uint32_t *__restrict__ table, int ntiles_cap) {
  __shared__ uint32_t h[256];
  const int n = *d_n;
  const int ntiles = (n + RS_TILE - 1) / RS_TILE;
  if ((int)blockIdx.x >= ntiles)
    return;
  h[threadIdx.x] = 0;
  __syncthreads();
  const int base = blockIdx.x * RS_TILE;
  for (int k = 0; k < RS_ITEMS; k++) {
    int i = base + k * RS_BLOCK + threadIdx.x;
    if (i < n)
      atomicAdd(&h[(keys[i] >> shift) & 255u], 1u);
  }
  __syncthreads();
  table[threadIdx.x * ntiles_cap + blockIdx.x] = h[threadIdx.x];
}

// Row scan of the digit-major table: CTA d scans row d (the counts of digit d over the live tiles) in place to exclusive
// prefixes and leaves the digit total in dtot[d]. 256 CTAs, all loads in flight at once; the scatter kernel turns the 256
// digit totals into digit bases itself (one block scan), so there is no serial pass over the table.
__global__ void __launch_bounds__(256) k_radix_scan(uint32_t *__restrict__ table, const int *__restrict__ d_n, int ntiles_cap,
                                                   uint32_t *__restrict__ dtot) {
  const int n = *d_n;
  const int ntiles = (n + RS_TILE - 1) / RS_TILE;
  uint32_t *row = table + (size_t)blockIdx.x * ntiles_cap;
  uint32_t carry = 0;
  for (int base = 0; base < ntiles; base += 256) {
    const int t = base + threadIdx.x;
    const uint32_t v = t < ntiles ? row[t] : 0u;
    uint32_t tot;
    const uint32_t ex = block_excl_scan<256>(v, &tot);
    if (t < ntiles)
      row[t] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0)
    dtot[blockIdx.x] = carry;
}

__global__ void __launch_bounds__(RS_BLOCK) k_radix_scatter(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals,
                                                           const int *__restrict__ d_n, int shift,
                                                           const uint32_t *__restrict__ table, int ntiles_cap,
                                                           const uint32_t *__restrict__ dtot, uint32_t *__restrict__ keys_out,
                                                           uint32_t *__restrict__ vals_out) {
  __shared__ uint32_t wcnt[RS_WARPS][256]; // per-warp digit counts -> per-warp bases
  const int n = *d_n;
  const int ntiles = (n + RS_TILE - 1) / RS_TILE;
  if ((int)blockIdx.x >= ntiles)
    return;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_BLOCK)
    (&wcnt[0][0])[i] = 0;
  __syncthreads();
  // warp w owns the contiguous chunk [base + w*CH, base + (w+1)*CH), processed 32 at a time in order
  constexpr int CH = RS_TILE / RS_WARPS; // 512
  const int wbase = blockIdx.x * RS_TILE + wid * CH;
  uint32_t myk[CH / 32], myv[CH / 32], myrank[CH / 32];
#pragma unroll
  for (int r = 0; r < CH / 32; r++) {
    int i = wbase + r * 32 + lane;
    bool ok = i < n;
    uint32_t k = ok ? keys[i] : 0xffffffffu;
    myk[r] = k;
    myv[r] = ok ? vals[i] : 0;
    uint32_t d = (k >> shift) & 255u;
    unsigned act = __ballot_sync(0xffffffffu, ok);
    uint32_t rank = 0;
    if (ok) {
      unsigned peers = __match_any_sync(act, d);
      rank = wcnt[wid][d] + __popc(peers & ((1u << lane) - 1u));
      __syncwarp(act);
      if ((peers & ((1u << lane) - 1u)) == 0)
        wcnt[wid][d] += __popc(peers);
    }
    __syncwarp();
    myrank[r] = rank;
  }
  __syncthreads();
  // per digit: exclusive prefix over warps + global base of this tile
  {
    const int d = threadIdx.x; // 256 threads == 256 digits
    const uint32_t dbase = block_excl_scan<RS_BLOCK>(dtot[d], nullptr); // elements with a smaller digit
    uint32_t run = dbase + table[d * ntiles_cap + blockIdx.x];
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) {
      uint32_t c = wcnt[w][d];
      wcnt[w][d] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < CH / 32; r++) {
    int i = wbase + r * 32 + lane;
    if (i < n) {
      uint32_t d = (myk[r] >> shift) & 255u;
      uint32_t pos = wcnt[wid][d] + myrank[r];
      keys_out[pos] = myk[r];
      vals_out[pos] = myv[r];
    }
  }
}

} // namespace ml
