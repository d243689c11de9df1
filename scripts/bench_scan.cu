// Micro-benchmark behind the choice of the exclusive-scan kernel used for the kNN grid (13.5 M cell counters on the G-64 map)
// and for the point-sized scans: thread-contiguous vs warp-striped tiles, single-pass look-back vs reduce+apply.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I mins_b200/csrc scripts/bench_scan.cu -o build_ab/bench_scan
#include "prims.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
namespace ml {
// pass 1 of the two-pass scan: tile sums; the CTA that finishes last turns them into exclusive tile prefixes in place
// (tile_sums[active] = total). `done` must be 0 on entry and is 0 again on exit.
template <int ITEMS>
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_reduce(uint32_t *in, const int *__restrict__ d_n, uint32_t *tile_sums,
                                                           unsigned int *done) {
  constexpr int TILE = SCAN_BLOCK * ITEMS;
  __shared__ uint32_t s_last;
  const int n = *d_n;
  const uint32_t active = n > 0 ? (uint32_t)((n + TILE - 1) / TILE) : 1u;
  if (blockIdx.x >= active)
    return;
  ScanTile<ITEMS, false> t;
  t.load(in, (long long)blockIdx.x * TILE, n);
  const uint32_t tot = t.reduce();
  if (threadIdx.x == 0) {
    *(volatile uint32_t *)&tile_sums[blockIdx.x] = tot;
    __threadfence();
    s_last = atomicAdd(done, 1u) == active - 1u;
  }
  __syncthreads();
  if (!s_last)
    return;
  __threadfence();
  // one CTA, thread t owns a contiguous chunk of the tile sums
  const uint32_t per = (active + SCAN_BLOCK - 1) / SCAN_BLOCK;
  const uint32_t lo = min(threadIdx.x * per, active), hi = min(lo + per, active);
  volatile uint32_t *ts = tile_sums;
  uint32_t s = 0;
  for (uint32_t i = lo; i < hi; i++)
    s += ts[i];
  uint32_t total;
  uint32_t e = block_excl_scan<SCAN_BLOCK>(s, &total);
  for (uint32_t i = lo; i < hi; i++) {
    const uint32_t v = ts[i];
    ts[i] = e;
    e += v;
  }
  if (threadIdx.x == 0) {
    ts[active] = total;
    *done = 0u;
  }
}

// pass 2: scan inside the tile on top of the tile's exclusive prefix
template <int ITEMS, bool ZERO_IN>
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_apply(uint32_t *in, const int *__restrict__ d_n, const uint32_t *__restrict__ tile_sums,
                                                          uint32_t *__restrict__ out, int *__restrict__ d_total) {
  constexpr int TILE = SCAN_BLOCK * ITEMS;
  const int n = *d_n;
  const uint32_t active = n > 0 ? (uint32_t)((n + TILE - 1) / TILE) : 1u;
  if (blockIdx.x >= active)
    return;
  const long long base = (long long)blockIdx.x * TILE;
  const uint32_t prefix = tile_sums[blockIdx.x];
  ScanTile<ITEMS, ZERO_IN> t;
  t.load(in, base, n);
  const uint32_t tot = t.scan();
  t.store(out, base, n, prefix);
  if (base + TILE >= (long long)n && threadIdx.x == 0) {
    out[n] = prefix + tot;
    if (d_total)
      *d_total = (int)(prefix + tot);
  }
}

} // namespace ml
using namespace ml;

#define CK(x)                                                                                                          \
  do {                                                                                                                 \
    cudaError_t e = (x);                                                                                               \
    if (e != cudaSuccess) {                                                                                            \
      printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__);                                                \
      exit(1);                                                                                                         \
    }                                                                                                                  \
  } while (0)

__global__ void k_flush(float4 *p, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    p[i] = make_float4(1.f, 2.f, 3.f, 4.f);
}

template <typename F> float time_it(F &&f, int reps, bool flush, float4 *fl, size_t fln, cudaStream_t s) {
  cudaEvent_t a, b;
  cudaEventCreate(&a), cudaEventCreate(&b);
  float tot = 0;
  for (int r = 0; r < reps + 3; r++) {
    if (flush)
      k_flush<<<148 * 8, 256, 0, s>>>(fl, fln);
    cudaEventRecord(a, s);
    f();
    cudaEventRecord(b, s);
    CK(cudaStreamSynchronize(s));
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    if (r >= 3)
      tot += ms;
  }
  return tot / reps * 1e3f;
}

int main() {
  cudaStream_t s;
  CK(cudaStreamCreate(&s));
  const size_t fln = (256u << 20) / 16;
  float4 *fl;
  CK(cudaMalloc(&fl, fln * 16));
  unsigned long long *state;
  unsigned int *ctr, *done;
  uint32_t *tsum;
  CK(cudaMalloc(&state, 8 * 70000));
  CK(cudaMemset(state, 0, 8 * 70000));
  CK(cudaMalloc(&ctr, 4));
  CK(cudaMemset(ctr, 0, 4));
  CK(cudaMalloc(&done, 4));
  CK(cudaMemset(done, 0, 4));
  CK(cudaMalloc(&tsum, 4 * 70000));
  for (int n : {13500000, 1131072, 131072, 1000}) {
    std::vector<uint32_t> hin(n), ref(n + 1), got(n + 1);
    uint32_t x = 12345u;
    for (int i = 0; i < n; i++) {
      x = x * 1664525u + 1013904223u;
      hin[i] = ((x >> 16) % 14 == 0) ? 1 + ((x >> 8) & 3) : 0;
    }
    ref[0] = 0;
    for (int i = 0; i < n; i++)
      ref[i + 1] = ref[i] + hin[i];
    uint32_t *in, *out;
    int *dn;
    CK(cudaMalloc(&in, 4 * (size_t)(n + 8)));
    CK(cudaMalloc(&out, 4 * (size_t)(n + 8)));
    CK(cudaMalloc(&dn, 4));
    CK(cudaMemcpy(in, hin.data(), 4 * (size_t)n, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dn, &n, 4, cudaMemcpyHostToDevice));
    uint32_t epoch = 1;
    auto check = [&](const char *name, float us_warm, float us_cold) {
      CK(cudaMemcpy(got.data(), out, 4 * (size_t)(n + 1), cudaMemcpyDeviceToHost));
      bool ok = got == ref;
      printf("n=%9d %-34s warm %7.1f us  L2-flushed %7.1f us  %s\n", n, name, us_warm, us_cold, ok ? "ok" : "MISMATCH");
      CK(cudaMemset(out, 0xff, 4 * (size_t)(n + 1)));
    };
#define RUN(name, body)                                                                                                \
  {                                                                                                                    \
    auto f = [&]() { body; };                                                                                          \
    float w = time_it(f, 20, false, fl, fln, s), c = time_it(f, 20, true, fl, fln, s);                                 \
    check(name, w, c);                                                                                                 \
  }
    RUN("lookback striped ITEMS=8", (k_scan_lb<8, false><<<(n + 2047) / 2048 + 3, SCAN_BLOCK, 0, s>>>(in, dn, out, nullptr, state, ctr, epoch++)));
    RUN("lookback striped ITEMS=16", (k_scan_lb<16, false><<<(n + 4095) / 4096 + 3, SCAN_BLOCK, 0, s>>>(in, dn, out, nullptr, state, ctr, epoch++)));
    RUN("lookback striped ITEMS=32", (k_scan_lb<32, false><<<(n + 8191) / 8192 + 3, SCAN_BLOCK, 0, s>>>(in, dn, out, nullptr, state, ctr, epoch++)));
    RUN("reduce+apply striped ITEMS=8", (k_scan_reduce<8><<<(n + 2047) / 2048 + 3, SCAN_BLOCK, 0, s>>>(in, dn, tsum, done),
                                         k_scan_apply<8, false><<<(n + 2047) / 2048 + 3, SCAN_BLOCK, 0, s>>>(in, dn, tsum, out, nullptr)));
    RUN("reduce+apply striped ITEMS=16", (k_scan_reduce<16><<<(n + 4095) / 4096 + 3, SCAN_BLOCK, 0, s>>>(in, dn, tsum, done),
                                          k_scan_apply<16, false><<<(n + 4095) / 4096 + 3, SCAN_BLOCK, 0, s>>>(in, dn, tsum, out, nullptr)));
    RUN("reduce+apply striped ITEMS=32", (k_scan_reduce<32><<<(n + 8191) / 8192 + 3, SCAN_BLOCK, 0, s>>>(in, dn, tsum, done),
                                          k_scan_apply<32, false><<<(n + 8191) / 8192 + 3, SCAN_BLOCK, 0, s>>>(in, dn, tsum, out, nullptr)));
    cudaFree(in), cudaFree(out), cudaFree(dn);
  }
  return 0;
}
