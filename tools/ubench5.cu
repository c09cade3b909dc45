// Prototype of the "q form" MLSA sample step (mlsa_seg.cu, round 2): the stage output is one FMA behind its
// input, the exchange overlaps the tap chain. What does one scheduler of a B200 make of it, by warps per
// scheduler, with and without the LOP3 ties that pin ptxas's interleaving?
//   TIE  0: no tie (ptxas hoists the chain)  1: chain step j tied to the dot product of tap j - DIST
//   EXCH 0: u = y * 1e-3 (no exchange)       1: STS / __syncwarp / 13 LDS / Pade combine
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o tools/ubench5 tools/ubench5.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kT = 34;
__device__ __forceinline__ double tie(double x, double y, int zero) {
  return __hiloint2double(__double2hiint(x) ^ (__double2hiint(y) & zero), __double2loint(x));
}

template <int TIE, int DIST, int EXCH, int WPS>
__global__ void __launch_bounds__(128, WPS) k_q(double* out, const double* in, int nsamples, long long* cyc, int zero) {
  __shared__ __align__(16) double X[4][2][128];
  __shared__ __align__(16) double In[4][8 * 64];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane < 30 ? lane / 5 : 5, s = lane < 30 ? lane - 5 * g : lane - 30;
  const int g5 = 5 * g, prev_lane = s > 0 ? g5 + s - 1 : g5;
  const double a = in[0], ma = -a, iaa = 1.0 - a * a, ppw = in[1 + s];
  double Q[kT], W[kT], winc[kT];
#pragma unroll
  for (int k = 0; k < kT; ++k) { Q[k] = 0.0; W[k] = in[8 + k] * (1 + 0.01 * s); winc[k] = in[48 + k] * 1e-6; }
  for (int k = lane; k < 8 * 64; k += 32) In[wib][k] = 1.0 + 1e-3 * k;
  for (int k = lane; k < 256; k += 32) X[wib][0][k] = 0.0;
  __syncwarp();
  double ad11 = 0, gin = 0, uin = 0, Y0 = 0, K = 0.1, c0 = W[0];
  const long long t0 = clock64();
  for (int n = 0; n < nsamples; ++n) {
    double* Xp = X[wib][n & 1];
    const double y = __fma_rn(K, uin, Y0);
    const double d11 = __fma_rn(iaa, gin, ad11);
    const double fo = d11 * c0;
    double wy1 = 0, wy2 = 0, wy3 = 0, wy4 = 0, wy5 = 0, wf1 = 0, wf2 = 0, wf3 = 0, wf4 = 0, wf5 = 0, yprev = 0, fprev = 0, e = 1.0;
    if (EXCH) {
      Xp[lane] = y; Xp[32 + lane] = fo; Xp[64 + lane] = ppw * y; Xp[96 + lane] = ppw * fo;
      __syncwarp();
      const double* Wx = Xp + 64 + g5;
      wy1 = Wx[0]; wy2 = Wx[1]; wy3 = Wx[2]; wy4 = Wx[3]; wy5 = Wx[4];
      wf1 = Wx[32]; wf2 = Wx[33]; wf3 = Wx[34]; wf4 = Wx[35]; wf5 = Wx[36];
      yprev = Xp[prev_lane]; fprev = Xp[32 + prev_lane];
      e = In[wib][g * 64 + (n & 63)];
    }
    double pj2 = 0.0, pj1 = __fma_rn(iaa, uin, Q[0]);
    ad11 = a * d11;
    W[1] += winc[1];
    const double v = ma * W[1];
    double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
    for (int j = 1; j <= kT + 1; ++j) {
      double pj = 0.0;
      if (j < kT) {
        double qj = Q[j];
        if (TIE && j >= DIST + 2) qj = tie(qj, ((j - DIST) & 1) ? acc1 : acc0, zero);
        pj = __fma_rn(ma, pj1, qj);
      }
      if (j == 1) Q[0] = a * pj1;
      if (j >= 2 && j - 1 < kT) Q[j - 1] = __fma_rn(a, pj1, pj2);
      if (j == 2) acc0 = v * Q[0];
      if (j >= 3) { if ((j - 2) & 1) acc1 = __fma_rn(W[j - 2], Q[j - 2], acc1); else acc0 = __fma_rn(W[j - 2], Q[j - 2], acc0); }
      if (j + 1 < kT) W[j + 1] += winc[j + 1];
      pj2 = pj1; pj1 = pj;
    }
    W[0] += winc[0];
    if (EXCH) {
      const double Af = wf5 + (wf3 + wf1), Bf = wf4 + wf2, Ay = wy5 + (wy3 + wy1), By = wy4 + wy2;
      const double d120 = (e + Af) - Bf, x1 = __fma_rn(2.0, Af, e), d220 = (x1 + Ay) - By;
      gin = s == 0 ? d120 : fprev;
      uin = s == 0 ? d220 : yprev;
      uin = uin * 1e-3 + 1.0;  // keep the recursion bounded whatever the random coefficients are
    } else {
      gin = fo * 1e-3;
      uin = y * 1e-3 + 1.0;
    }
    Y0 = (acc0 + acc1) * 1e-3;
    K = v * iaa;
    c0 = W[0];
  }
  const long long t1 = clock64();
  double acc = Y0 + K + ad11 + gin + uin;
#pragma unroll
  for (int k = 0; k < kT; ++k) acc += Q[k] + W[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

static double h_in[128];
static double *d_out, *d_in;
static long long* d_cyc;

template <int TIE, int DIST, int EXCH, int WPS>
void run(const char* name) {
  const int ns = 12000;
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, k_q<TIE, DIST, EXCH, WPS>);
  k_q<TIE, DIST, EXCH, WPS><<<148 * WPS, 128>>>(d_out, d_in, 200, d_cyc, 0);
  cudaDeviceSynchronize();
  k_q<TIE, DIST, EXCH, WPS><<<148 * WPS, 128>>>(d_out, d_in, ns, d_cyc, 0);
  cudaDeviceSynchronize();
  static long long hc[1024];
  cudaMemcpy(hc, d_cyc, sizeof(long long) * 148 * WPS, cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < 148 * WPS; ++i) mx = hc[i] > mx ? hc[i] : mx;
  cudaError_t e = cudaGetLastError();
  printf("%-28s TIE%d DIST%d X%d WPS%d regs=%3d spill=%zu  cyc/sample=%7.1f  per warp-sample=%6.1f %s\n", name, TIE, DIST, EXCH,
         WPS, fa.numRegs, fa.localSizeBytes, (double)mx / ns, (double)mx / ns / WPS, e == cudaSuccess ? "" : cudaGetErrorString(e));
  fflush(stdout);
}

int main() {
  cudaMalloc(&d_out, (size_t)148 * 8 * 1024 * 8);
  cudaMalloc(&d_in, sizeof h_in);
  cudaMalloc(&d_cyc, 2048 * 8);
  for (int i = 0; i < 128; ++i) h_in[i] = 0.01 * ((i * 7919) % 97) / 97.0;
  h_in[0] = 0.55; h_in[1] = 0.4999391; h_in[2] = 0.1107098; h_in[3] = 0.01369984; h_in[4] = 0.0009564853; h_in[5] = 0.00003041721;
  cudaMemcpy(d_in, h_in, sizeof h_in, cudaMemcpyHostToDevice);
  k_q<0, 4, 0, 2><<<296, 128>>>(d_out, d_in, 300000, d_cyc, 0);  // clocks up
  cudaDeviceSynchronize();
  run<0, 4, 0, 1>("no tie, no exchange");
  run<0, 4, 0, 2>("no tie, no exchange");
  run<1, 4, 0, 1>("tie 4, no exchange");
  run<1, 4, 0, 2>("tie 4, no exchange");
  run<1, 2, 0, 1>("tie 2, no exchange");
  run<1, 2, 0, 2>("tie 2, no exchange");
  run<1, 8, 0, 1>("tie 8, no exchange");
  run<1, 8, 0, 2>("tie 8, no exchange");
  run<0, 4, 1, 1>("no tie, exchange");
  run<0, 4, 1, 2>("no tie, exchange");
  run<1, 4, 1, 1>("tie 4, exchange");
  run<1, 4, 1, 2>("tie 4, exchange");
  run<1, 8, 1, 1>("tie 8, exchange");
  run<1, 8, 1, 2>("tie 8, exchange");
  run<1, 2, 1, 1>("tie 2, exchange");
  run<1, 2, 1, 2>("tie 2, exchange");
  return 0;
}
