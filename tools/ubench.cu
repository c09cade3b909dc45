// Micro-benchmarks that size the MLSA kernel design on B200 (sm_100a): FP64 latency /
// throughput per SMSP as a function of warps and ILP, SHFL latency, LDS latency.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench tools/ubench.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void k_dfma(double* out, int iters, double a, double b, long long* cyc) {
  double x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = threadIdx.x + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = __fma_rn(x[i], a, b);
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

__global__ void k_shfl_chain(double* out, int iters, long long* cyc) {
  double x = threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) x = __shfl_xor_sync(0xffffffffu, x, 1) + 1.0;
  long long t1 = clock64();
  out[threadIdx.x] = x;
  if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_dadd_chain(double* out, int iters, long long* cyc) {
  double x = threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) x = x + 1.0;
  long long t1 = clock64();
  out[threadIdx.x] = x;
  if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_lds_chain(double* out, int iters, long long* cyc) {
  __shared__ double s[64];
  s[threadIdx.x] = (double)((threadIdx.x + 1) & 31);
  s[threadIdx.x + 32] = 0;
  __syncthreads();
  int idx = threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) idx = (int)s[idx];
  long long t1 = clock64();
  out[threadIdx.x] = idx;
  if (threadIdx.x == 0) *cyc = t1 - t0;
}

template <int ILP>
void run_dfma(int blocks_per_sm, int threads) {
  int sms;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out;
  long long* cyc;
  cudaMalloc(&out, (size_t)sms * blocks_per_sm * threads * 8);
  cudaMalloc(&cyc, 8);
  const int iters = 20000;
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  k_dfma<ILP><<<sms * blocks_per_sm, threads>>>(out, iters, 0.9999, 1e-9, cyc);
  cudaEventRecord(a);
  k_dfma<ILP><<<sms * blocks_per_sm, threads>>>(out, iters, 0.9999, 1e-9, cyc);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  long long c;
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  double fma = (double)iters * ILP * sms * blocks_per_sm * threads;
  printf("dfma ILP=%d warps/SM=%d: %.2f TFLOP/s, %.2f cyc/iter (=%.2f cyc per dep FMA step), %.1f lane-FMA/clk/SM\n", ILP,
         blocks_per_sm * threads / 32, 2 * fma / (ms * 1e-3) / 1e12, (double)c / iters, (double)c / iters,
         fma / ((double)c * sms));
  cudaFree(out);
  cudaFree(cyc);
}

int main() {
  double* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 20);
  cudaMalloc(&cyc, 8);
  long long c;
  const int iters = 100000;
  k_dadd_chain<<<1, 32>>>(out, iters, cyc);
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("DADD dependent latency: %.2f cyc\n", (double)c / iters);
  k_shfl_chain<<<1, 32>>>(out, iters, cyc);
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("SHFL(64-bit)+DADD dependent latency: %.2f cyc\n", (double)c / iters);
  k_lds_chain<<<1, 32>>>(out, iters, cyc);
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("LDS.64 + cvt dependent latency: %.2f cyc\n", (double)c / iters);
  // single warp per SMSP (4 warps per SM), varying ILP
  run_dfma<1>(1, 128);
  run_dfma<2>(1, 128);
  run_dfma<4>(1, 128);
  run_dfma<8>(1, 128);
  run_dfma<16>(1, 128);
  // ILP 1, varying warps per SM
  run_dfma<1>(1, 256);
  run_dfma<1>(1, 512);
  run_dfma<1>(1, 1024);
  run_dfma<1>(2, 1024);
  run_dfma<4>(1, 256);
  run_dfma<4>(1, 512);
  run_dfma<8>(2, 1024);
  return 0;
}
