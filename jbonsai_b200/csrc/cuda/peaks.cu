// Roofline denominators measured live on the device the bench runs on:
//   FP64 FMA peak (DFMA micro-kernel, 8 independent chains per thread) — the MLSA kernel's bound
//   HBM copy bandwidth (read + write bytes of a plain copy)          — the MLPG kernel's bound
// Measurement helpers only; nothing on the synthesis path calls them.
#include <cuda_runtime.h>

#include "kernels.h"

namespace jbk {

namespace {

__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; ++i) {
    x0 = __fma_rn(x0, a, b); x1 = __fma_rn(x1, a, b); x2 = __fma_rn(x2, a, b); x3 = __fma_rn(x3, a, b);
    x4 = __fma_rn(x4, a, b); x5 = __fma_rn(x5, a, b); x6 = __fma_rn(x6, a, b); x7 = __fma_rn(x7, a, b);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

__global__ void k_copy(const double2* __restrict__ src, double2* __restrict__ dst, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = src[i];
}

}  // namespace

// Returns 0 or a cudaError_t. tflops: FP64 FMA rate (2 flop per FMA); gbs: copy read+write GB/s.
int measure_peaks(double* tflops, double* gbs) {
  cudaError_t e;
  int dev = 0, sms = 0;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return (int)e;
  if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return (int)e;
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  double* buf = nullptr;
  const int blocks = sms * 8, threads = 256, iters = 1 << 16;
  if ((e = cudaMalloc((void**)&buf, (size_t)blocks * threads * 8)) != cudaSuccess) return (int)e;
  float best = 1e30f;
  for (int r = 0; r < 4; ++r) {
    cudaEventRecord(a);
    k_dfma<<<blocks, threads>>>(buf, iters, 0.999999, 1e-9);
    cudaEventRecord(b);
    if ((e = cudaEventSynchronize(b)) != cudaSuccess) return (int)e;
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    if (r > 0 && ms < best) best = ms;
  }
  *tflops = 2.0 * 8.0 * iters * (double)blocks * threads / (best * 1e-3) / 1e12;
  cudaFree(buf);

  const size_t n = (size_t)1 << 27;  // 2 GiB per buffer as double2
  double2 *s = nullptr, *d = nullptr;
  if ((e = cudaMalloc((void**)&s, n * 16)) != cudaSuccess) return (int)e;
  if ((e = cudaMalloc((void**)&d, n * 16)) != cudaSuccess) { cudaFree(s); return (int)e; }
  cudaMemset(s, 0, n * 16);
  best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(a);
    k_copy<<<sms * 16, 512>>>(s, d, n);
    cudaEventRecord(b);
    if ((e = cudaEventSynchronize(b)) != cudaSuccess) return (int)e;
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    if (r > 0 && ms < best) best = ms;
  }
  *gbs = 2.0 * n * 16 / (best * 1e-3) / 1e9;
  cudaFree(s);
  cudaFree(d);
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  return (int)cudaGetLastError();
}

}  // namespace jbk
