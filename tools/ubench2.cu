// SHFL throughput and smem-exchange latency (sizes the cross-lane steps of the MLSA kernel).
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_shfl_tput(double* out, int iters, long long* cyc) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    x0 = __shfl_xor_sync(~0u, x0, 1); x1 = __shfl_xor_sync(~0u, x1, 2); x2 = __shfl_xor_sync(~0u, x2, 4); x3 = __shfl_xor_sync(~0u, x3, 8);
    x4 = __shfl_xor_sync(~0u, x4, 1); x5 = __shfl_xor_sync(~0u, x5, 2); x6 = __shfl_xor_sync(~0u, x6, 4); x7 = __shfl_xor_sync(~0u, x7, 8);
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

// dependent chain: STS.64 -> syncwarp -> LDS.64 (neighbour's value) -> DADD
__global__ void k_smem_xchg(double* out, int iters, long long* cyc) {
  __shared__ double s[128];
  double x = threadIdx.x;
  double* my = s + threadIdx.x;
  const double* nb = s + (threadIdx.x ^ 1);
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    *(volatile double*)my = x;
    __syncwarp();
    x = *(volatile const double*)nb + 1.0;
    __syncwarp();
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

// mixed: DFMA stream (ILP 4) with S shuffles per 8 DFMA: does SHFL steal DP issue?
template <int S>
__global__ void k_mix(double* out, int iters, double a, double b, long long* cyc) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3;
  double s0 = x0, s1 = x1, s2 = x2, s3 = x3;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    x0 = __fma_rn(x0, a, b); x1 = __fma_rn(x1, a, b); x2 = __fma_rn(x2, a, b); x3 = __fma_rn(x3, a, b);
    if (S > 0) s0 = __shfl_xor_sync(~0u, s0, 1);
    if (S > 1) s1 = __shfl_xor_sync(~0u, s1, 2);
    x0 = __fma_rn(x0, a, b); x1 = __fma_rn(x1, a, b); x2 = __fma_rn(x2, a, b); x3 = __fma_rn(x3, a, b);
    if (S > 2) s2 = __shfl_xor_sync(~0u, s2, 4);
    if (S > 3) s3 = __shfl_xor_sync(~0u, s3, 8);
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + s0 + s1 + s2 + s3;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
  double* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 24);
  cudaMalloc(&cyc, 8);
  long long c;
  const int iters = 20000;
  for (int warps = 1; warps <= 8; warps *= 2) {
    k_shfl_tput<<<148, 128 * warps>>>(out, iters, cyc);
    cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("SHFL.64 x8 independent, %d warps/SMSP: %.2f cyc per SHFL.64 per warp, %.2f cyc per SHFL.64 per SMSP\n", warps,
           (double)c / iters / 8, (double)c / iters / 8 / warps);
  }
  k_smem_xchg<<<1, 32>>>(out, iters, cyc);
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("STS.64 -> syncwarp -> LDS.64 -> DADD -> syncwarp chain: %.2f cyc\n", (double)c / iters);
  for (int warps = 1; warps <= 2; warps *= 2) {
#define MIX(S)                                                                                   \
  k_mix<S><<<148, 128 * warps>>>(out, iters, 0.9999, 1e-9, cyc);                                   \
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);                                                \
  printf("8 DFMA + %d SHFL.64 per iter, %d warps/SMSP: %.2f cyc/iter/warp\n", S, warps, (double)c / iters);
    MIX(0) MIX(1) MIX(2) MIX(4)
  }
  return 0;
}
