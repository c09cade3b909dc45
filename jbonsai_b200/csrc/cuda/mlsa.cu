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
// Schedule: a lone warp issues in order, so the loop-carried chains are interleaved by hand and
// every shared-memory round trip is covered with independent arithmetic:
//   region 1  K = KE + kappa u ; partial y = yz + K cp  (yz = sum c_i z_i and cp = sum c_i (-a)^(i+1)
//             are formed off the critical path) -> publish ; t' = z + P K
//   sync 1
//   region 2  loads in flight: q/z scan of the NEXT sample from t', coefficient interpolation;
//             then lanes of stage s+1 add the 6 partials of stage s (their next input; stage 1
//             adds stage 5) ; df1 section one sample ahead ; publish y, f and E
//   sync 2
//   region 3  loads in flight: yz of the next sample; then ONE combine for both cascades (lanes 0-17:
//             d22[0]/output of sample n from the five y's; lanes 18-31: d12[0]/x of sample n+1 from
//             the five df1 outputs) and the next carry KE from the E's.
// The loop is unrolled by two so that the two exchange buffers are compile-time addresses; the last
// sample is peeled (no look-ahead beyond the launch). Cross-lane traffic is shared memory (an
// LDS.128 broadcast is one wavefront; SHFL.64 costs 8 issue cycles per SMSP on B200). Stores are
// unconditional (per-lane dump slots): no divergent branch (predicating the traffic down to the lanes
// that need it was measured: fewer wavefronts, but more registers/instructions and no gain). The input ring is filled by cp.async.
#include <type_traits>

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
  const int wG = R ? oG : oDump + 2 * lane;
  const int rGin = (s >= 1 && s < 5) ? oV + 8 + s - 1 : oG;  // d12[s] of the current sample

  for (int i = lane; i < 2 * kBufD; i += 32) sBuf[i] = 0.0;
  for (int i = lane; i < kCH * 32; i += 32) sOut[i] = 0.0;

  // ---- state (layout: kernels.h)
  double t[6] = {0, 0, 0, 0, 0, 0}, tm1 = 0.0, c[6] = {0, 0, 0, 0, 0, 0}, c1 = 0.0;
  double d11 = 0.0, gin = 0.0, uin = 0.0;
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
    gin = stp[176 + s];   // d12[s]: input of section s+1
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
    if (f0 == 0) {  // is_first: coefficients = mc2b(first frame), before the post-filter
      const double* b1 = d.b_first ? d.b_first + (int64_t)u * kNM : bf;
#pragma unroll
      for (int i = 0; i < 6; ++i) c[i] = cvalid[i] ? b1[jtap[i]] : 0.0;
      c1 = b1[1];
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) cinc[i] = (cc[i] - c[i]) / fper;
    c1inc = (cc1 - c1) / fper;
  }
  int left = fp, fnext = f0 + 2;  // samples left in the frame; frame whose coefficients to prefetch next

  double* wv = d.wave + (foff + f0) * (int64_t)fp;
  sIn[lane] = lane < N ? wv[lane] : 0.0;  // chunks 0 and 1
  __syncwarp();

  double* const B0 = sBuf;
  double* const B1 = sBuf + kBufD;
  const double* const vpE = R ? B1 + oV + 8 : B0 + oV;  // region-3 operands of even / odd samples
  const double* const vpO = R ? B0 + oV + 8 : B1 + oV;

  // ---- prologue: bring the pipeline to the top of sample 0 of this launch
  //   z, KE: zero-carry scan and carry of sample 0; yz, cp: dot-product pieces; buffer 0 holds
  //   f_1..f_5 and (d12[0], x after df1) of sample 0
  double z[6], KE, yz, cp, cpinc;
  {
    d11 = __fma_rn(a, d11, iaa * gin);  // df1 section s+1 for sample 0 (mlsa.rs:57-58)
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
    B0[wVf] = fk;
    __syncwarp();
    const double2 e01 = *reinterpret_cast<const double2*>(B1 + rE);
    const double2 e23 = *reinterpret_cast<const double2*>(B1 + rE + 2);
    const double e4 = B1[rE + 4];
    KE = __fma_rn(W[4], e4, __fma_rn(W[2], e23.x, W[0] * e01.x)) + __fma_rn(W[3], e23.y, W[1] * e01.y);
    const double2 f12 = *reinterpret_cast<const double2*>(B0 + oV + 8);
    const double2 f34 = *reinterpret_cast<const double2*>(B0 + oV + 10);
    const double f5 = B0[oV + 12];
    const double e0 = sIn[0];
    const double A = __fma_rn(pp5, f5, __fma_rn(pp3, f34.x, pp1 * f12.x));
    const double Bv = __fma_rn(pp4, f34.y, pp2 * f12.y);
    B0[oG] = (e0 + A) - Bv;  // every lane writes the same pair
    B0[oG + 1] = __fma_rn(2.0, A, e0);
    __syncwarp();
  }
  auto dot6 = [](const double (&x)[6], const double (&y)[6]) {
    return __fma_rn(x[4], y[4], __fma_rn(x[2], y[2], x[0] * y[0])) +
           __fma_rn(x[5], y[5], __fma_rn(x[3], y[3], x[1] * y[1]));
  };
  yz = dot6(c, z);
  cp = dot6(c, P);
  cpinc = dot6(cinc, P);

  int inIdx = 1;                 // ring slot of the next sample's excitation
  double* outp = sOut + lane;    // this lane's column of the output tile

  auto step = [&](double* __restrict__ B, double* __restrict__ Bn, const double* __restrict__ vp, auto LAST) {
    constexpr bool last = decltype(LAST)::value;
    // ---------------- region 1: critical path  u -> K -> partial y
    const double K = __fma_rn(kap, uin, KE);
    B[wYP] = __fma_rn(K, cp, yz);
    t[0] = __fma_rn(P[0], K, z[0]);
    t[1] = __fma_rn(P[1], K, z[1]);
    t[2] = __fma_rn(P[2], K, z[2]);
    t[3] = __fma_rn(P[3], K, z[3]);
    t[4] = __fma_rn(P[4], K, z[4]);
    t[5] = __fma_rn(P[5], K, z[5]);
    tm1 = bb > 0 ? K : 0.0;
    __syncwarp();
    // ---------------- region 2
    const double2 r01 = *reinterpret_cast<const double2*>(B + rYP);  // 6 partials of stage `instage`
    const double2 r23 = *reinterpret_cast<const double2*>(B + rYP + 2);
    const double2 r45 = *reinterpret_cast<const double2*>(B + rYP + 4);
    gin = B[rGin];                                                     // d12[s] of this sample
    const double base = *(R ? sIn + inIdx : B + oG + 1);               // R = 1: e(n+1);  R = 0: x after df1 (n)
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
    cp += cpinc;
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
      cp = dot6(c, P);
      cpinc = dot6(cinc, P);
      const double* bn = bsrc + (int64_t)(fnext < f1 ? fnext : f1 - 1) * kNM;
#pragma unroll
      for (int i = 0; i < 6; ++i) ccn[i] = cvalid[i] ? bn[jtap[i]] : 0.0;
      ccn1 = bn[1];
      ++fnext;
      left = fp;
    }
    const double ysum = ((r01.x + r01.y) + (r23.x + r23.y)) + (r45.x + r45.y);
    B[wVy] = ysum;
    if (!last) {  // df1 section s+1, one sample ahead (mlsa.rs:57-58)
      d11 = __fma_rn(a, d11, iaa * gin);
      Bn[wVf] = d11 * c1;
    }
    __syncwarp();
    // ---------------- region 3
    const double2 v12 = *reinterpret_cast<const double2*>(vp);
    const double2 v34 = *reinterpret_cast<const double2*>(vp + 2);
    const double v5 = vp[4];
    const double2 e01 = *reinterpret_cast<const double2*>(B + rE);
    const double2 e23 = *reinterpret_cast<const double2*>(B + rE + 2);
    const double e4 = B[rE + 4];
    yz = dot6(c, z);
    const double A = __fma_rn(pp5, v5, __fma_rn(pp3, v34.x, pp1 * v12.x));
    const double Bv = __fma_rn(pp4, v34.y, pp2 * v12.y);
    const double r1 = (base + A) - Bv;         // R = 0: d22[0](n)    R = 1: d12[0](n+1)
    const double r2 = __fma_rn(2.0, A, base);  // R = 0: output(n)    R = 1: x after df1 (n+1)
    KE = __fma_rn(W[4], e4, __fma_rn(W[2], e23.x, W[0] * e01.x)) + __fma_rn(W[3], e23.y, W[1] * e01.y);
    uin = s == 0 ? r1 : ysum;
    *outp = r2;
    outp += 32;
    if (!last) *reinterpret_cast<double2*>(Bn + wG) = make_double2(r1, r2);
    inIdx = (inIdx + 1) & (2 * kCH - 1);
  };

  const std::false_type F{};
  const std::true_type L{};
  for (int i0 = 0; i0 < N; i0 += kCH) {
    outp = sOut + lane;
#pragma unroll 1
    for (int j = 0; j < kCH - 2; j += 2) {
      step(B0, B1, vpE, F);
      step(B1, B0, vpO, F);
    }
    cp_async_wait_all();  // the next chunk has landed; the sync inside the next step publishes it
    step(B0, B1, vpE, F);
    if (i0 + kCH >= N) step(B1, B0, vpO, L); else step(B1, B0, vpO, F);
    __syncwarp();
    if (lane < kCH) {  // flush 16 outputs, refill this half of the ring
      wv[i0 + lane] = sOut[lane * 32] * p.volume;
      if (i0 + 2 * kCH + lane < N) cp_async8(&sIn[(i0 & kCH) + lane], wv + i0 + 2 * kCH + lane);
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
        stp[176 + s] = gin;  // d12[s]
        stp[182 + s] = uin;  // d22[s]
      }
    }
    if (lane == 0) {
      stp[188 + 1] = c1;
      stp[188] = bsrc[(int64_t)(f1 - 1) * kNM];  // coefficients[0] = cc[0] at a frame end
      stp[176 + 5] = B1[oV + 8 + 4];             // d12[5], d22[5]: inputs of no stage, kept for layout parity
      stp[182 + 5] = B1[oV + 4];
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
