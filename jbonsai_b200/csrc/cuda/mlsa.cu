// MLSA filter, variant 2 (default): one warp per utterance, software-pipelined blocked scan.
//
// Reference: MelLogSpectrumApproximation::<6>::{df, df1, df2, fir} (src/vocoder/mlsa.rs:48-163)
// driven per sample by Vocoder::synthesize (src/vocoder/mod.rs:112-141).
//
// Layout: lane = 6*s + b owns taps 6b+1 .. 6b+6 of the warped delay line d21[s] of Pade stage
// s+1 (s = 0..4, b = 0..5; taps 35/36 and lanes 30/31 are padding with zero coefficients).
//
// Arithmetic (same filter, re-associated; parity bar = 1e-6 abs on the waveform, tests):
//   * all-pass chain in "q form": t'_j = (t_{j-1} + a t_j) - a t'_{j-1}  (== iaa*rem_j + a*d_j of
//     mlsa.rs:141-153): a first-order scan over the tap index with ratio -a;
//   * blocked scan: each lane scans its 6 taps with zero carry-in (z_i); block aggregates
//     E_b = z_5 are exchanged; the carry into block b is K_b = sum_{m<b} (-a)^(6(b-1-m)) E_m
//     + kappa_b u (u = stage input d22[s] of the previous sample) and t'_i = z_i + (-a)^(i+1) K_b;
//   * df1/df2 sign pattern (mlsa.rs:54-79): with A = sum over odd stages of pp_k v_k and B = sum
//     over even stages, d[0] = (x + A) - B and x' = x + 2A.
//
// Schedule: a lone warp issues in order, so the two loop-carried chains are interleaved by hand:
//   region 1  K, t' (6 FMA), dot product with the coefficients -> partial y;  q/z scan of the NEXT
//             sample from the new t';  coefficient interpolation;  publish partial y and E
//   sync 1
//   region 2  lanes of stage s+1 sum the 6 partials of stage s (their next input; stage 1 sums
//             stage 5) ; next carry K^E from the E's ; df1 section one sample ahead ; publish
//   sync 2
//   region 3  ONE combine for both cascades: lanes 0-17 form d22[0]/output of sample n from the
//             five y's, lanes 18-31 form d12[0]/x of sample n+1 from the five df1 outputs.
// Cross-lane traffic is double-buffered shared memory (an LDS.128 broadcast is one wavefront;
// SHFL.64 costs 8 issue cycles per SMSP on B200). Stores are unconditional (per-lane dump
// slots) so the loop has no divergent branch. The input chunk ring is filled with cp.async.
#include "kernels.h"

namespace jbk {

namespace {

__constant__ double c_pp[6] = {1.00000000000, 0.49993910000, 0.11070980000,
                               0.01369984000, 0.00095648530, 0.00003041721};  // mlsa.rs:25-32

__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gmem_src) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(sa), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.wait_all;\n" ::: "memory");
}

constexpr int kNM = 35;   // mel-cepstral order + 1
constexpr int kCH = 16;   // samples per staged chunk (ring of 2 chunks)

// Exchange buffer (doubles), two of them alternate by sample parity:
//   [0..40)   E   block aggregates, by lane          [40..80)  YP  partial dot products, by lane
//   [80..85)  y_1..y_5   [88..93) f_1..f_5 (df1 outputs, one sample ahead)
//   [96..98)  (d12[0], x after df1) of the next sample      [100..164) per-lane dump slots (2 each)
constexpr int kBufD = 164;
constexpr int oE = 0, oYP = 40, oV = 80, oG = 96, oDump = 100;

__global__ void __launch_bounds__(32) k_mlsa_v2(BatchLayout b, VocoderParams p, VocoderDev d) {
  __shared__ __align__(16) double sBuf[2 * kBufD];
  __shared__ __align__(16) double sIn[2 * kCH];
  __shared__ __align__(16) double sOut[kCH * 32];

  const int lane = threadIdx.x;
  const int u = b.order ? b.order[blockIdx.x] : (int)blockIdx.x;
  const int64_t foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  int f0 = d.frame_begin, f1 = f0 + d.frame_count;
  if (f1 > T || f1 < f0) f1 = T;
  if (f0 >= f1) return;
  const int fp = p.fperiod;
  const int N = (f1 - f0) * fp;  // samples of this launch; fp % kCH == 0 (launcher)

  const int s = lane / 6, bb = lane - 6 * s;   // s == 5: padding lanes
  const bool real = s < 5, pub = real && bb == 0;
  const bool R = lane >= 18;                   // combine role: 0 = df2 (sample n), 1 = df1 (sample n+1)
  const int instage = (s >= 1 && s < 5) ? s - 1 : 4;  // group whose partial sums this lane adds up
  const double a = p.alpha, ma = -a, iaa = 1.0 - a * a;
  double P[6], W[5], kap;
  {
    double pw = ma;
#pragma unroll
    for (int i = 0; i < 6; ++i) { P[i] = pw; pw *= ma; }   // (-a)^(i+1)
    const double A6 = P[5];
#pragma unroll
    for (int m = 0; m < 5; ++m) {
      double w = 0.0;
      if (m < bb) { w = 1.0; for (int e = 0; e < bb - 1 - m; ++e) w *= A6; }
      W[m] = w;                                             // (-a)^(6(b-1-m)), m < b
    }
    pw = 1.0;
    for (int e = 0; e < 6 * bb; ++e) pw *= ma;
    kap = real ? iaa * pw / ma : 0.0;                       // iaa (-a)^(6b-1); b = 0: -iaa/a
  }
  const double pp1 = c_pp[1], pp2 = c_pp[2], pp3 = c_pp[3], pp4 = c_pp[4], pp5 = c_pp[5];
  const double fper = (double)fp;

  // per-lane offsets into an exchange buffer
  const int wE = oE + lane, wYP = oYP + lane;
  const int rE = oE + 6 * s, rYP = oYP + 6 * instage;
  const int wVy = pub ? oV + instage : oDump + 2 * lane;
  const int wVf = pub ? oV + 8 + s : oDump + 2 * lane;
  const int rV = oV + (R ? 8 : 0);
  const int rVf = oV + 8 + ((s >= 1 && s < 5) ? s - 1 : 0);
  const int wG = R ? oG : oDump + 2 * lane;

  for (int i = lane; i < 2 * kBufD; i += 32) sBuf[i] = 0.0;
  for (int i = lane; i < kCH * 32; i += 32) sOut[i] = 0.0;

  // ---- state (layout: kernels.h)
  double t[6] = {0, 0, 0, 0, 0, 0}, tm1 = 0.0, c[6] = {0, 0, 0, 0, 0, 0}, c1 = 0.0;
  double d11 = 0.0, ginf = 0.0, uin = 0.0;
  int jtap[6];
  bool cvalid[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    jtap[i] = 6 * bb + 1 + i;
    cvalid[i] = real && jtap[i] >= 2 && jtap[i] < kNM;
  }
  double* stp = d.state ? d.state + (int64_t)u * kVocStateDoubles : nullptr;
  const bool resume = stp != nullptr && f0 > 0;
  if (resume && real) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      t[i] = jtap[i] < kNM ? stp[s * (kNM - 1) + jtap[i] - 1] : 0.0;
      c[i] = cvalid[i] ? stp[188 + jtap[i]] : 0.0;
    }
    tm1 = bb > 0 ? stp[226 + 5 * s + bb - 1] : 0.0;
    c1 = stp[188 + 1];
    d11 = stp[170 + s + 1];
    ginf = stp[176 + s];  // d12[s]: input of section s+1 (s = 0: d12[0])
    uin = stp[182 + s];   // d22[s]: input of stage s+1
  }

  // ---- coefficients of the first frame (mod.rs:79-125); the next frame's are prefetched
  const double* bsrc = d.b + foff * kNM;
  double cc[6], ccn[6], cinc[6], cc1, ccn1, c1inc;
  {
    const double* bf = bsrc + (int64_t)f0 * kNM;
    const double* bn = bsrc + (int64_t)(f0 + 1 < f1 ? f0 + 1 : f0) * kNM;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      cc[i] = cvalid[i] ? bf[jtap[i]] : 0.0;
      ccn[i] = cvalid[i] ? bn[jtap[i]] : 0.0;
    }
    cc1 = bf[1];
    ccn1 = bn[1];
    if (f0 == 0) {  // is_first: coefficients = mc2b(first frame)
#pragma unroll
      for (int i = 0; i < 6; ++i) c[i] = cc[i];
      c1 = cc1;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) cinc[i] = (cc[i] - c[i]) / fper;
    c1inc = (cc1 - c1) / fper;
  }
  int left = fp, fnext = f0 + 2;  // samples left in the frame; frame whose coefficients to prefetch next

  double* wv = d.wave + (foff + f0) * (int64_t)fp;
  sIn[lane] = lane < N ? wv[lane] : 0.0;  // chunks 0 and 1
  __syncwarp();

  // ---- prologue: bring the pipeline to the top of sample 0 of this launch
  //   z, KE: zero-carry scan and carry of sample 0;  (g0n, x1n): df1 result of sample 0;  ginf: d12[s] of sample 0
  double z[6], KE, g0n, x1n;
  {
    double* B1 = sBuf + kBufD;
    d11 = __fma_rn(a, d11, iaa * ginf);  // df1 section s+1 for sample 0 (mlsa.rs:57-58)
    const double fk = d11 * c1;
    const double q0 = __fma_rn(a, t[0], tm1), q1 = __fma_rn(a, t[1], t[0]), q2 = __fma_rn(a, t[2], t[1]),
                 q3 = __fma_rn(a, t[3], t[2]), q4 = __fma_rn(a, t[4], t[3]), q5 = __fma_rn(a, t[5], t[4]);
    z[0] = q0;
    z[1] = __fma_rn(ma, z[0], q1);
    z[2] = __fma_rn(ma, z[1], q2);
    z[3] = __fma_rn(ma, z[2], q3);
    z[4] = __fma_rn(ma, z[3], q4);
    z[5] = __fma_rn(ma, z[4], q5);
    B1[wE] = z[5];
    B1[wVf] = fk;
    __syncwarp();
    const double2 e01 = *reinterpret_cast<const double2*>(B1 + rE);
    const double2 e23 = *reinterpret_cast<const double2*>(B1 + rE + 2);
    const double e4 = B1[rE + 4];
    KE = __fma_rn(W[4], e4, __fma_rn(W[2], e23.x, W[0] * e01.x)) + __fma_rn(W[3], e23.y, W[1] * e01.y);
    const double2 f12 = *reinterpret_cast<const double2*>(B1 + oV + 8);
    const double2 f34 = *reinterpret_cast<const double2*>(B1 + oV + 10);
    const double f5 = B1[oV + 12];
    const double e0 = sIn[0];
    const double A = __fma_rn(pp5, f5, __fma_rn(pp3, f34.x, pp1 * f12.x));
    const double Bv = __fma_rn(pp4, f34.y, pp2 * f12.y);
    g0n = (e0 + A) - Bv;
    x1n = __fma_rn(2.0, A, e0);
    if (s >= 1) ginf = B1[rVf];  // d12[s](0) = f_s(0)
    __syncwarp();
    sBuf[oG] = g0n;  // every lane holds the same pair
    sBuf[oG + 1] = x1n;
  }

  int po = 0;  // offset of the current exchange buffer (0 or kBufD)
#pragma unroll 1
  for (int n = 0; n < N; ++n) {
    double* B = sBuf + po;
    const bool has_next = n + 1 < N;
    // ---------------- region 1
    const double K = __fma_rn(kap, uin, KE);
    t[0] = __fma_rn(P[0], K, z[0]);
    t[1] = __fma_rn(P[1], K, z[1]);
    t[2] = __fma_rn(P[2], K, z[2]);
    t[3] = __fma_rn(P[3], K, z[3]);
    t[4] = __fma_rn(P[4], K, z[4]);
    t[5] = __fma_rn(P[5], K, z[5]);
    tm1 = bb > 0 ? K : 0.0;
    const double ya = __fma_rn(c[4], t[4], __fma_rn(c[2], t[2], c[0] * t[0]));
    const double yb = __fma_rn(c[5], t[5], __fma_rn(c[3], t[3], c[1] * t[1]));
    B[wYP] = ya + yb;
    {  // zero-carry scan of the next sample
      const double q0 = __fma_rn(a, t[0], tm1), q1 = __fma_rn(a, t[1], t[0]), q2 = __fma_rn(a, t[2], t[1]),
                   q3 = __fma_rn(a, t[3], t[2]), q4 = __fma_rn(a, t[4], t[3]), q5 = __fma_rn(a, t[5], t[4]);
      z[0] = q0;
      z[1] = __fma_rn(ma, z[0], q1);
      z[2] = __fma_rn(ma, z[1], q2);
      z[3] = __fma_rn(ma, z[2], q3);
      z[4] = __fma_rn(ma, z[3], q4);
      z[5] = __fma_rn(ma, z[4], q5);
    }
    B[wE] = z[5];
#pragma unroll
    for (int i = 0; i < 6; ++i) c[i] += cinc[i];  // mod.rs:133-135
    c1 += c1inc;
    if (--left == 0) {  // frame end: *coefficients = cc (mod.rs:140), then the next frame's prologue
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        c[i] = cc[i];
        cc[i] = ccn[i];
        cinc[i] = (cc[i] - c[i]) / fper;
      }
      c1 = cc1;
      cc1 = ccn1;
      c1inc = (cc1 - c1) / fper;
      const double* bn = bsrc + (int64_t)(fnext < f1 ? fnext : f1 - 1) * kNM;
#pragma unroll
      for (int i = 0; i < 6; ++i) ccn[i] = cvalid[i] ? bn[jtap[i]] : 0.0;
      ccn1 = bn[1];
      ++fnext;
      left = fp;
    }
    if ((n & (kCH - 1)) == kCH - 2) cp_async_wait_all();  // the chunk after this one has landed
    __syncwarp();
    // ---------------- region 2
    const double2 r01 = *reinterpret_cast<const double2*>(B + rYP);
    const double2 r23 = *reinterpret_cast<const double2*>(B + rYP + 2);
    const double2 r45 = *reinterpret_cast<const double2*>(B + rYP + 4);
    const double2 e01 = *reinterpret_cast<const double2*>(B + rE);
    const double2 e23 = *reinterpret_cast<const double2*>(B + rE + 2);
    const double e4 = B[rE + 4];
    const double2 gx = *reinterpret_cast<const double2*>(B + oG);  // (d12[0], x after df1) of sample n
    const double e_next = sIn[(n + 1) & (2 * kCH - 1)];
    const double ysum = ((r01.x + r01.y) + (r23.x + r23.y)) + (r45.x + r45.y);
    KE = __fma_rn(W[4], e4, __fma_rn(W[2], e23.x, W[0] * e01.x)) + __fma_rn(W[3], e23.y, W[1] * e01.y);
    g0n = gx.x;
    x1n = gx.y;
    // df1 section s+1, one sample ahead (mlsa.rs:57-58)
    const double d11n = __fma_rn(a, d11, iaa * (s == 0 ? g0n : ginf));
    const double fk = d11n * c1;
    d11 = has_next ? d11n : d11;
    B[wVy] = ysum;
    B[wVf] = fk;
    __syncwarp();
    // ---------------- region 3: one combine for both cascades (mlsa.rs:59-65, 72-78)
    const double2 v12 = *reinterpret_cast<const double2*>(B + rV);
    const double2 v34 = *reinterpret_cast<const double2*>(B + rV + 2);
    const double v5 = B[rV + 4];
    const double gf = B[rVf];
    const double base = R ? e_next : x1n;
    const double A = __fma_rn(pp5, v5, __fma_rn(pp3, v34.x, pp1 * v12.x));
    const double Bv = __fma_rn(pp4, v34.y, pp2 * v12.y);
    const double r1 = (base + A) - Bv;      // R = 0: d22[0](n)    R = 1: d12[0](n+1)
    const double r2 = __fma_rn(2.0, A, base);  // R = 0: output(n)    R = 1: x after df1 (n+1)
    uin = s == 0 ? r1 : ysum;
    ginf = (has_next && s >= 1) ? gf : ginf;
    sOut[(n & (kCH - 1)) * 32 + lane] = r2;
    po ^= kBufD;
    {
      double* Bn = sBuf + po;
      Bn[wG] = r1;
      Bn[wG + 1] = r2;
    }
    if ((n & (kCH - 1)) == kCH - 1) {  // chunk done: flush 16 outputs, refill this half of the ring
      __syncwarp();
      const int i0 = n - (kCH - 1);
      if (lane < kCH) {
        wv[i0 + lane] = sOut[lane * 32] * p.volume;
        if (i0 + 2 * kCH + lane < N) cp_async8(&sIn[(i0 & kCH) + lane], wv + i0 + 2 * kCH + lane);
      }
    }
  }

  // ---- streaming state (same layout as variants 0/1, kernels.h)
  if (stp) {
    __syncwarp();
    if (real) {
#pragma unroll
      for (int i = 0; i < 6; ++i)
        if (jtap[i] < kNM) {
          stp[s * (kNM - 1) + jtap[i] - 1] = t[i];
          if (s == 0 && cvalid[i]) stp[188 + jtap[i]] = c[i];
        }
      if (bb > 0) stp[226 + 5 * s + bb - 1] = tm1;
      if (bb == 0) {
        stp[170 + s + 1] = d11;
        stp[176 + s] = s == 0 ? g0n : ginf;  // d12[s]
        stp[182 + s] = uin;                  // d22[s]
      }
    }
    if (lane == 0) {
      stp[188 + 1] = c1;
      stp[188] = bsrc[(int64_t)(f1 - 1) * kNM];  // coefficients[0] = cc[0] at a frame end
      stp[176 + 5] = sBuf[po + oV + 8 + 4];      // d12[5], d22[5]: inputs of no stage, kept for layout parity
      stp[182 + 5] = sBuf[(po ^ kBufD) + oV + 4];
    }
  }
}

}  // namespace

int launch_mlsa_v2(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, cudaStream_t stream) {
  if (b.n_utt == 0) return 0;
  if (p.nmcp != kNM || p.fperiod % kCH != 0 || !(p.alpha > 1e-3)) return (int)cudaErrorInvalidValue;
  k_mlsa_v2<<<b.n_utt, 32, 0, stream>>>(b, p, d);
  return (int)cudaGetLastError();
}

}  // namespace jbk
