// FP64 issue-rate probe: how many cycles one warp-wide FP64 instruction costs a scheduler, by
// operand pattern. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o ubench3 ubench3.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(double* out, const double* in, int iters) {
  double acc[8], x[8], z[8];
  for (int j = 0; j < 8; ++j) { acc[j] = in[j]; x[j] = in[8 + j + threadIdx.x % 3]; z[j] = in[16 + j + threadIdx.x % 5]; }
  const double c1 = in[30], c2 = in[31];
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (MODE == 0) acc[j] = __fma_rn(acc[j], c1, c2);              // 1 varying operand
        if (MODE == 1) acc[j] = __fma_rn(x[j], c1, acc[j]);            // 2 varying
        if (MODE == 2) acc[j] = __fma_rn(x[j], z[(j + r) & 7], acc[j]);  // 3 varying
        if (MODE == 3) acc[j] = __dadd_rn(acc[j], x[j]);               // DADD 2 varying
        if (MODE == 4) acc[j] = __dmul_rn(acc[j], c1);                 // DMUL
        if (MODE == 5) { x[j] = __fma_rn(c1, x[j], z[j]); acc[j] = __fma_rn(z[(j + 1) & 7], x[(j + 3) & 7], acc[j]); }
      }
    }
  }
  long long t1 = clock64();
  double s = 0;
  for (int j = 0; j < 8; ++j) s += acc[j] + x[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (double)(t1 - t0);
}

template <int MODE>
void run(const char* name, int warps_per_smsp, double* out, double* in) {
  const int iters = 4000;
  const int per_iter = (MODE == 5 ? 64 : 32);
  k<MODE><<<148, 128 * warps_per_smsp>>>(out, in, iters);
  cudaDeviceSynchronize();
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  k<MODE><<<148, 128 * warps_per_smsp>>>(out, in, iters);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double cyc; cudaMemcpy(&cyc, out, 8, cudaMemcpyDeviceToHost);
  const double instr = (double)iters * per_iter * warps_per_smsp;  // warp instructions per scheduler
  printf("%-34s warps/smsp=%d  cycles/warp-instr(per scheduler)=%.3f  (%.1f ms)\n", name, warps_per_smsp, cyc / instr, ms);
}

int main() {
  double *out, *in;
  cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&in, 64 * 8);
  double h[64]; for (int i = 0; i < 64; ++i) h[i] = 1.0 + 1e-9 * i;
  cudaMemcpy(in, h, sizeof h, cudaMemcpyHostToDevice);
  for (int w = 1; w <= 4; w *= 2) {
    run<0>("DFMA acc=fma(acc,c,c)", w, out, in);
    run<1>("DFMA acc=fma(x,c,acc)", w, out, in);
    run<2>("DFMA acc=fma(x,z,acc) 3 regs", w, out, in);
    run<3>("DADD acc=acc+x", w, out, in);
    run<4>("DMUL acc=acc*c", w, out, in);
    run<5>("pair: x=fma(c,x,z); acc=fma(z,x,acc)", w, out, in);
  }
  return 0;
}
