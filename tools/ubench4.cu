// Inner-loop prototypes of the MLSA sample step (round 2): what bounds one scheduler of a B200 when a lane
// runs the 34-tap q-form chain of mlsa_seg.cu, by design variant and by resident warps per scheduler.
//   CMODE 0: c[34], cinc[34] in registers, 34 DADD per sample (round-1 layout)
//   CMODE 1: c shared by the five stage lanes of a slot through shared memory: each lane advances 7 of the 34
//            coefficients and stores them, every lane reads all of them back as broadcast LDS.128
//   CMODE 2: constant coefficients (no interpolation) - the floor of chain + dot product
//   LA   1 : two-tap look-ahead (chain of 17 instead of 34, one more FMA per tap pair)
//   EXCH 1 : the per-sample exchange (STS, __syncwarp, 12 LDS, Pade combine) as in k_mlsa_v3; 0: u = y * 1e-3
//   NCH    : independent delay lines per lane (same coefficients)
// Also: dependent-DFMA latency, LDS.64/LDS.128 broadcast wavefronts, SHFL rate.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o tools/ubench4 tools/ubench4.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kT = 34;

template <int CMODE, int LA, int EXCH, int NCH, int WPS>
__global__ void __launch_bounds__(128, WPS) k_step(double* out, const double* in, int nsamples, long long* cyc) {
  __shared__ __align__(16) double Y[4][2][32];
  __shared__ __align__(16) double F[4][2][32];
  __shared__ __align__(16) double C[4][2][18 * 16];  // [warp][parity][tap pair][slot 0..7][2]
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane < 30 ? lane / 5 : 5, s = lane < 30 ? lane - 5 * g : lane - 30;
  const int g5 = 5 * g;
  const int prev_lane = s > 0 ? g5 + s - 1 : g5;
  const double a = in[0], ma = -a, iaa = 1.0 - a * a, aa = a * a;
  const double pp1 = in[1], pp2 = in[2], pp3 = in[3], pp4 = in[4], pp5 = in[5];
  double t[NCH][kT], c[kT], cinc[kT];
#pragma unroll
  for (int k = 0; k < kT; ++k) {
    c[k] = in[8 + k] * (1 + 0.01 * s);
    cinc[k] = in[48 + k] * 1e-6;
#pragma unroll
    for (int h = 0; h < NCH; ++h) t[h][k] = 0.0;
  }
  // CMODE 1: this lane advances coefficients 7 s .. 7 s + 6 (lane s == 4: 6 of them) of slot g
  double oc[7], oci[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    const int kk = 7 * s + k < kT ? 7 * s + k : kT - 1;
    oc[k] = in[8 + kk];
    oci[k] = in[48 + kk] * 1e-6;
  }
  if (CMODE == 1) {
    for (int k = lane; k < 18 * 16; k += 32) { C[wib][0][k] = 0.01; C[wib][1][k] = 0.01; }
  }
  Y[wib][0][lane] = 0; Y[wib][1][lane] = 0; F[wib][0][lane] = 0; F[wib][1][lane] = 0;
  __syncwarp();
  double uin[NCH], d11 = 0, gin = 0;
#pragma unroll
  for (int h = 0; h < NCH; ++h) uin[h] = 0.0;
  double e = 1.0;
  const long long t0 = clock64();
  for (int n = 0; n < nsamples; ++n) {
    const int par = n & 1;
    double ysum[NCH];
    const double2* Cr = reinterpret_cast<const double2*>(&C[wib][par][0]) + g;
#pragma unroll
    for (int h = 0; h < NCH; ++h) {
      double po = t[h][0];
      double pn = __fma_rn(a, po, iaa * uin[h]);
      t[h][0] = pn;
      double y0 = 0.0, y1 = 0.0;
      if (LA == 0) {
#pragma unroll
        for (int i = 1; i < kT; i += 2) {
          double c0, c1;
          if (CMODE == 1) { const double2 cc = Cr[(i >> 1) * 8]; c0 = cc.x; c1 = cc.y; }
          else { c0 = c[i]; c1 = i + 1 < kT ? c[i + 1] : 0.0; }
#ifdef RU
          // dot product of a tap issued right behind the chain step that reads the same tap value in the same
          // operand slot (operand reuse cache)
          double q = __fma_rn(a, t[h][i], po);
          po = t[h][i];
          const double p1 = __fma_rn(pn, ma, q);
          t[h][i] = p1;
          if (i + 1 < kT) {
            q = __fma_rn(a, t[h][i + 1], po);
            po = t[h][i + 1];
            pn = __fma_rn(p1, ma, q);
            y0 = __fma_rn(p1, c0, y0);
            t[h][i + 1] = pn;
            y1 = __fma_rn(pn, c1, y1);
          } else {
            pn = p1;
            y0 = __fma_rn(p1, c0, y0);
          }
#else
          double q = __fma_rn(a, t[h][i], po);
          po = t[h][i];
          pn = __fma_rn(ma, pn, q);
          t[h][i] = pn;
          y0 = __fma_rn(c0, pn, y0);
          if (i + 1 < kT) {
            q = __fma_rn(a, t[h][i + 1], po);
            po = t[h][i + 1];
            pn = __fma_rn(ma, pn, q);
            t[h][i + 1] = pn;
            y1 = __fma_rn(c1, pn, y1);
          }
#endif
        }
      } else {
#pragma unroll
        for (int i = 1; i < kT; i += 2) {
          double c0, c1;
          if (CMODE == 1) { const double2 cc = Cr[(i >> 1) * 8]; c0 = cc.x; c1 = cc.y; }
          else { c0 = c[i]; c1 = i + 1 < kT ? c[i + 1] : 0.0; }
          const double q0 = __fma_rn(a, t[h][i], po);
          po = t[h][i];
          if (i + 1 < kT) {
            const double q1 = __fma_rn(a, t[h][i + 1], po);
            po = t[h][i + 1];
            const double r = __fma_rn(ma, q0, q1);
            const double p0 = __fma_rn(ma, pn, q0);   // off the chain
            pn = __fma_rn(aa, pn, r);                 // chain: two taps per step
            t[h][i] = p0;
            t[h][i + 1] = pn;
            y0 = __fma_rn(c0, p0, y0);
            y1 = __fma_rn(c1, pn, y1);
          } else {
            pn = __fma_rn(ma, pn, q0);
            t[h][i] = pn;
            y0 = __fma_rn(c0, pn, y0);
          }
        }
      }
      ysum[h] = y0 + y1;
    }
    if (CMODE == 0) {
#pragma unroll
      for (int k = 0; k < kT; ++k) c[k] += cinc[k];
    } else if (CMODE == 1) {
      double* Cw = &C[wib][par ^ 1][0];
#pragma unroll
      for (int k = 0; k < 7; ++k) {
        oc[k] += oci[k];
        const int kk = 7 * s + k;  // tap index; pair kk>>1, element kk&1
        if (kk < kT && lane < 30) Cw[((kk >> 1) * 8 + g) * 2 + (kk & 1)] = oc[k];
      }
    }
    if (EXCH) {
      d11 = __fma_rn(a, d11, iaa * gin);
      Y[wib][par][lane] = ysum[0];
      F[wib][par][lane] = d11 * (CMODE == 1 ? oc[0] : c[0]);
      __syncwarp();
      const double* Yp = Y[wib][par];
      const double* Fp = F[wib][par];
      const double y1v = Yp[g5], y2v = Yp[g5 + 1], y3v = Yp[g5 + 2], y4v = Yp[g5 + 3], y5v = Yp[g5 + 4];
      const double f1v = Fp[g5], f2v = Fp[g5 + 1], f3v = Fp[g5 + 2], f4v = Fp[g5 + 3], f5v = Fp[g5 + 4];
      const double yprev = Yp[prev_lane], fprev = Fp[prev_lane];
      const double Af = __fma_rn(pp5, f5v, __fma_rn(pp3, f3v, pp1 * f1v));
      const double Bf = __fma_rn(pp4, f4v, pp2 * f2v);
      const double Ay = __fma_rn(pp5, y5v, __fma_rn(pp3, y3v, pp1 * y1v));
      const double By = __fma_rn(pp4, y4v, pp2 * y2v);
      const double d120 = (e + Af) - Bf;
      const double x1 = __fma_rn(2.0, Af, e);
      const double d220 = (x1 + Ay) - By;
      e = __fma_rn(2.0, Ay, x1) * 1e-3 + 1.0;
      gin = s == 0 ? d120 : fprev;
      uin[0] = s == 0 ? d220 : yprev;
#pragma unroll
      for (int h = 1; h < NCH; ++h) uin[h] = ysum[h] * 1e-3 + uin[0];
    } else {
      if (CMODE == 1) __syncwarp();
#pragma unroll
      for (int h = 0; h < NCH; ++h) uin[h] = ysum[h] * 1e-3 + 1.0;
    }
  }
  const long long t1 = clock64();
  double acc = e + d11;
#pragma unroll
  for (int h = 0; h < NCH; ++h) {
    acc += uin[h];
#pragma unroll
    for (int k = 0; k < kT; ++k) acc += t[h][k];
  }
  if (CMODE != 1) {
#pragma unroll
    for (int k = 0; k < kT; ++k) acc += c[k];
  }
#pragma unroll
  for (int k = 0; k < 7; ++k) acc += oc[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    cyc[blockIdx.x] = t1 - t0;
    cyc[1024 + blockIdx.x] = (long long)smid;
  }
}

static double h_in[128];
static double *d_out, *d_in;
static long long* d_cyc;

template <int CMODE, int LA, int EXCH, int NCH, int WPS>
void run(const char* name) {
  const int ns = 12000;
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, k_step<CMODE, LA, EXCH, NCH, WPS>);
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_step<CMODE, LA, EXCH, NCH, WPS>, 128, 0);
  k_step<CMODE, LA, EXCH, NCH, WPS><<<148 * WPS, 128>>>(d_out, d_in, 200, d_cyc);
  cudaDeviceSynchronize();
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  k_step<CMODE, LA, EXCH, NCH, WPS><<<148 * WPS, 128>>>(d_out, d_in, ns, d_cyc);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  static long long hc[2048];
  cudaMemcpy(hc, d_cyc, sizeof hc, cudaMemcpyDeviceToHost);
  long long cyc = 0, cmin = 1ll << 60;
  int per_sm[256] = {0}, sm_max = 0;
  for (int i = 0; i < 148 * WPS; ++i) {
    cyc = hc[i] > cyc ? hc[i] : cyc;
    cmin = hc[i] < cmin ? hc[i] : cmin;
    sm_max = ++per_sm[hc[1024 + i] & 255] > sm_max ? per_sm[hc[1024 + i] & 255] : sm_max;
  }
  printf("[blocks/SM max %d, cyc min %.1f] ", sm_max, (double)cmin / ns);
  cudaError_t e = cudaGetLastError();
  // chain-samples per scheduler-cycle: WPS warps x NCH chains per `cyc / ns` cycles
  const double cps = (double)cyc / ns;
  printf("%-44s C%d LA%d X%d NCH%d WPS%d regs=%3d spill=%zu occ=%d  cyc/sample=%7.1f  cyc/(warp-chain-sample)=%6.1f  %.2f ms (%.0f MHz) %s\n",
         name, CMODE, LA, EXCH, NCH, WPS, fa.numRegs, fa.localSizeBytes, occ, cps, cps / (WPS * NCH), ms, (double)cyc / ms * 1e-3,
         e == cudaSuccess ? "" : cudaGetErrorString(e));
  fflush(stdout);
}

// ---- basic probes
template <int ILP, int DEP3>
__global__ void k_lat(double* out, const double* in, int iters, long long* cyc) {
  double x[ILP], y = in[1], z = in[2];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = in[i] + threadIdx.x;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < ILP; ++i) x[i] = DEP3 ? __fma_rn(x[i], y, z) : __fma_rn(x[i], 0.999, 0.001);
  }
  const long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int ILP, int DEP3>
void run_lat(int wps) {
  const int iters = 2000;
  k_lat<ILP, DEP3><<<148, 128 * wps>>>(d_out, d_in, iters, d_cyc);
  cudaDeviceSynchronize();
  long long cyc;
  cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
  printf("DFMA chain: ILP=%d regs3=%d warps/smsp=%d  cycles per dependent step=%.2f  cycles/warp-instr/scheduler=%.2f\n", ILP,
         DEP3, wps, (double)cyc / (iters * 8), (double)cyc / (iters * 8.0 * ILP * wps));
  fflush(stdout);
}

// LDS patterns: MODE 0 = LDS.64 distinct per lane, 1 = LDS.64 broadcast by slot (lane/5), 2 = LDS.128 distinct,
// 3 = LDS.128 broadcast by slot, 4 = LDS.128 all lanes one address, 5 = SHFL.32 x2 (a 64-bit value)
template <int MODE>
__global__ void k_lds(double* out, int iters, long long* cyc) {
  __shared__ __align__(16) double S[16][256];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = lane; i < 256; i += 32) S[wib][i] = i;
  __syncwarp();
  const int g = lane / 5;
  double acc0 = 0, acc1 = 0, v = lane;
  const double* base = S[wib];
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int o = ((it + r) & 3) * 64;
      if (MODE == 0) acc0 += base[o + lane];
      if (MODE == 1) acc0 += base[o + g];
      if (MODE == 2) { const double2 x = *reinterpret_cast<const double2*>(base + o + 2 * lane); acc0 += x.x; acc1 += x.y; }
      if (MODE == 3) { const double2 x = *reinterpret_cast<const double2*>(base + o + 2 * g); acc0 += x.x; acc1 += x.y; }
      if (MODE == 4) { const double2 x = *reinterpret_cast<const double2*>(base + o); acc0 += x.x; acc1 += x.y; }
      if (MODE == 5) { v = __shfl_sync(0xffffffffu, v, (lane + r) & 31); acc0 += v; }
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc0 + acc1 + v;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE>
void run_lds(const char* name, int wps) {
  const int iters = 2000;
  k_lds<MODE><<<148, 128 * wps>>>(d_out, iters, d_cyc);
  cudaDeviceSynchronize();
  long long cyc;
  cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-34s warps/smsp=%d  SM cycles per warp-load=%.2f\n", name, wps, (double)cyc / (iters * 16.0 * wps * 4));
  { cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) printf("  error: %s\n", cudaGetErrorString(e)); }
  fflush(stdout);
}

int main() {
  cudaMalloc(&d_out, (size_t)148 * 8 * 1024 * 8);
  cudaMalloc(&d_in, sizeof h_in);
  cudaMalloc(&d_cyc, 2048 * 8);
  for (int i = 0; i < 128; ++i) h_in[i] = 0.01 * ((i * 7919) % 97) / 97.0;
  h_in[0] = 0.55; h_in[1] = 0.4999391; h_in[2] = 0.1107098; h_in[3] = 0.01369984; h_in[4] = 0.0009564853; h_in[5] = 0.00003041721;
  cudaMemcpy(d_in, h_in, sizeof h_in, cudaMemcpyHostToDevice);

  k_step<2, 0, 0, 1, 2><<<296, 128>>>(d_out, d_in, 400000, d_cyc);  // ~100 ms: clocks up
  cudaDeviceSynchronize();
  for (int w = 1; w <= 4; w *= 2) { run_lat<1, 0>(w); run_lat<1, 1>(w); run_lat<2, 1>(w); run_lat<4, 1>(w); run_lat<8, 1>(w); }
  for (int w = 1; w <= 4; w *= 2) {
    run_lds<0>("LDS.64 distinct", w); run_lds<1>("LDS.64 broadcast by slot", w); run_lds<2>("LDS.128 distinct", w);
    run_lds<3>("LDS.128 broadcast by slot", w); run_lds<4>("LDS.128 one address", w); run_lds<5>("SHFL.32 (x1)", w);
  }
  // the sample step, without and with the exchange
  run<2, 0, 0, 1, 1>("floor: chain+dot, c const");
  run<2, 0, 0, 1, 2>("floor: chain+dot, c const");
  run<2, 0, 0, 1, 3>("floor: chain+dot, c const");
  run<2, 0, 0, 1, 4>("floor: chain+dot, c const");
  run<0, 0, 0, 1, 1>("c in regs");
  run<0, 0, 0, 1, 2>("c in regs");
  run<1, 0, 0, 1, 1>("c shared via smem");
  run<1, 0, 0, 1, 2>("c shared via smem");
  run<1, 0, 0, 1, 3>("c shared via smem");
  run<1, 0, 0, 1, 4>("c shared via smem");
  run<0, 1, 0, 1, 1>("c in regs, look-ahead");
  run<0, 1, 0, 1, 2>("c in regs, look-ahead");
  run<1, 1, 0, 1, 1>("c shared, look-ahead");
  run<1, 1, 0, 1, 2>("c shared, look-ahead");
  run<1, 1, 0, 1, 3>("c shared, look-ahead");
  run<1, 0, 0, 2, 1>("c shared, 2 chains per lane");
  run<1, 0, 0, 2, 2>("c shared, 2 chains per lane");
  run<1, 1, 0, 2, 1>("c shared, 2 chains, look-ahead");
  run<1, 1, 0, 2, 2>("c shared, 2 chains, look-ahead");
  run<0, 0, 1, 1, 1>("c in regs + exchange (r1 kernel)");
  run<0, 0, 1, 1, 2>("c in regs + exchange (r1 kernel)");
  run<1, 0, 1, 1, 1>("c shared + exchange");
  run<1, 0, 1, 1, 2>("c shared + exchange");
  run<1, 0, 1, 1, 3>("c shared + exchange");
  run<1, 0, 1, 1, 4>("c shared + exchange");
  run<1, 1, 1, 1, 2>("c shared + exchange, look-ahead");
  run<1, 1, 1, 1, 3>("c shared + exchange, look-ahead");
  run<0, 1, 1, 1, 2>("c in regs + exchange, look-ahead");
  run<1, 0, 1, 2, 1>("c shared + exchange, 2 chains");
  run<1, 0, 1, 2, 2>("c shared + exchange, 2 chains");
  return 0;
}
