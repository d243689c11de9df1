// micro-benchmark: is the FP64 FMA rate the same on every SM, and does a co-resident neighbour SM of the TPC change it?
// one CTA per SM, 28 warps, a chain of dependent DFMAs per thread (8 independent chains); per-CTA %globaltimer duration by %smid.
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
#include <algorithm>
__global__ void __launch_bounds__(896, 1) k(double *out, unsigned long long *tl, int iters, int only_parity) {
  unsigned int smid;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  if (only_parity >= 0 && (int)(smid & 1) != only_parity) return;
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  double a[8];
  for (int q = 0; q < 8; q++) a[q] = 1.0 + threadIdx.x * 1e-9 + q;
  const double m = 1.0000001, c = 1e-9;
  for (int i = 0; i < iters; i++)
#pragma unroll
    for (int q = 0; q < 8; q++) a[q] = fma(a[q], m, c);
  double s = 0;
  for (int q = 0; q < 8; q++) s += a[q];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  __syncthreads();
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  if (threadIdx.x == 0) { tl[blockIdx.x * 2] = smid; tl[blockIdx.x * 2 + 1] = t1 - t0; }
}
int main() {
  int nsm = 148;
  double *out; unsigned long long *tl;
  cudaMalloc(&out, sizeof(double) * 148 * 896);
  cudaMalloc(&tl, sizeof(unsigned long long) * 148 * 2);
  for (int parity = -1; parity <= 1; parity++) {
    for (int rep = 0; rep < 3; rep++) {
      cudaMemset(tl, 0, sizeof(unsigned long long) * 148 * 2);
      k<<<nsm, 896>>>(out, tl, 2000, parity);
      cudaDeviceSynchronize();
    }
    std::vector<unsigned long long> h(148 * 2);
    cudaMemcpy(h.data(), tl, sizeof(unsigned long long) * 148 * 2, cudaMemcpyDeviceToHost);
    std::vector<std::pair<int, double>> v;
    for (int i = 0; i < 148; i++) if (h[2 * i + 1]) v.push_back({(int)h[2 * i], h[2 * i + 1] * 1e-3});
    std::sort(v.begin(), v.end());
    printf("parity %d: %zu CTAs ran; us by smid:", parity, v.size());
    for (auto &p : v) printf(" %d:%.1f", p.first, p.second);
    // 2000 iters * 8 DFMA * 28 warps = 448000 warp-DFMA per SM
    printf("\n");
  }
  return 0;
}
