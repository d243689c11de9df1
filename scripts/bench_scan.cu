// Micro-benchmark behind the choice of the exclusive-scan kernel used for the kNN grid (13.5 M cell counters on the G-64 map)
// and for the point-sized scans: thread-contiguous vs warp-striped tiles, single-pass look-back vs reduce+apply.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I mins_b200/csrc scripts/bench_scan.cu -o build_ab/bench_scan
#include "prims.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
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
