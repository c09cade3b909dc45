// Vocoder kernels (sm_100a), stage 0 (MLSA) path of Vocoder::synthesize
// (src/vocoder/mod.rs:72-141):
//   k_frame_prep  per frame : pitch period p (mod.rs:73-77) and b = mc2b(mcp) (cepstrum.rs:139-149)
//   k_pulse       per utt   : pitch counter / pulse train (excitation.rs:25-33,72-86,102-104)
//   k_excitation  per sample: noise + low-pass overlap-add in gather form (excitation.rs:43-64,
//                             67-100) and the gain exp(c[0]) (mod.rs:129-131)
//   k_mlsa_*      per utt   : Pade-cascaded MLSA filter df1/df2/fir (mlsa.rs:48-163) with the
//                             per-sample coefficient interpolation (mod.rs:118-123,133-140)
// Compile with -fmad=false; fused multiply-adds are written explicitly (__fma_rn) where the
// reference's FMA flavour of `fir` has them (mlsa.rs:82-115,147-153).
#include <algorithm>

#include "kernels.h"

namespace jbk {

namespace {

constexpr double kNoData = -1e10;
constexpr double kMaxLf0 = 9.903487552536127;
constexpr double kMinLf0 = 2.995732273553991;

__constant__ double c_ppade[6] = {1.00000000000, 0.49993910000, 0.11070980000,
                                  0.01369984000, 0.00095648530, 0.00003041721};

__device__ __forceinline__ int find_utt(const int64_t* frame_off, int n_utt, int64_t gf) {
  int lo = 0, hi = n_utt;  // largest u with frame_off[u] <= gf
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (frame_off[mid] <= gf) lo = mid; else hi = mid;
  }
  return lo;
}

// ---------------------------------------------------------------- frame prep
__global__ void k_frame_prep(VocoderParams p, VocoderDev d, int64_t total_frames) {
  int64_t gf = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gf >= total_frames) return;
  const double lf0 = d.lf0[gf];
  double pitch = 0.0;
  if (lf0 != kNoData) {
    double l = lf0 < kMinLf0 ? kMinLf0 : (lf0 > kMaxLf0 ? kMaxLf0 : lf0);
    pitch = p.rate / exp(l);
  }
  d.pitch[gf] = pitch;
  const double* c = d.mcp + gf * p.nmcp;
  double* b = d.b + gf * p.nmcp;
  const int last = p.nmcp - 1;
  double nxt = c[last];
  b[last] = nxt;
  for (int i = last - 1; i >= 0; --i) {
    nxt = (p.alpha != 0.0) ? c[i] - p.alpha * nxt : c[i];
    b[i] = nxt;
  }
}

// ---------------------------------------------------------------- pulse train
__global__ void k_pulse(BatchLayout b, VocoderParams p, VocoderDev d) {
  int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= b.n_utt) return;
  const int64_t foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  int f0 = d.frame_begin, f1 = f0 + d.frame_count;
  if (f1 > T || f1 < f0) f1 = T;
  double p_cur = 0.0, cnt = 0.0;
  if (d.state && f0 > 0) {
    p_cur = d.state[(int64_t)u * kVocStateDoubles + 223];
    cnt = d.state[(int64_t)u * kVocStateDoubles + 224];
  }
  const double fper = (double)p.fperiod;
  for (int f = f0; f < f1; ++f) {
    const double pitch = d.pitch[foff + f];
    double inc;
    if (p_cur != 0.0 && pitch != 0.0) {  // Excitation::start
      inc = (pitch - p_cur) / fper;
    } else {
      inc = 0.0;
      p_cur = pitch;
      cnt = pitch;
    }
    double* out = d.pulse + (foff + f) * p.fperiod;
    if (p_cur == 0.0) {
      for (int i = 0; i < p.fperiod; ++i) out[i] = 0.0;
    } else {
      for (int i = 0; i < p.fperiod; ++i) {
        cnt += 1.0;
        double pulse = 0.0;
        if (cnt >= p_cur) {
          cnt -= p_cur;
          pulse = sqrt(p_cur);
        }
        out[i] = pulse;
        p_cur += inc;
      }
    }
    p_cur = pitch;  // Excitation::end
  }
  if (d.state) {
    d.state[(int64_t)u * kVocStateDoubles + 223] = p_cur;
    d.state[(int64_t)u * kVocStateDoubles + 224] = cnt;
  }
}

// ---------------------------------------------------------------- excitation + gain
// One block per frame. Gather form of the ring-buffer overlap-add: output sample n sums, oldest
// source first, over source samples m = n-(nlpf-1) .. n:
//    [m == n-h]   + noise[m]                       (get_antipode_mut, h = (nlpf-1)/2)
//    [voiced(m)]  + (pulse[m]-noise[m]) * lpf_{frame(m)}[n-m]
// which is exactly the order in which the reference's scatter adds reach ring slot n.
constexpr int kMaxLpf = 64;
constexpr int kMaxFperiod = 1024;

__global__ void __launch_bounds__(256) k_excitation(BatchLayout b, VocoderParams p, VocoderDev d,
                                                    int64_t total_frames) {
  // block -> (utterance, frame) restricted to [frame_begin, frame_begin+frame_count)
  const int64_t gf = blockIdx.x;
  if (gf >= total_frames) return;
  const int u = find_utt(b.frame_off, b.n_utt, gf);
  const int64_t foff = b.frame_off[u];
  const int f = (int)(gf - foff);
  if (f < d.frame_begin || (d.frame_count >= 0 && f - d.frame_begin >= d.frame_count)) return;

  __shared__ double s_pulse[kMaxFperiod + kMaxLpf];
  __shared__ double s_noise[kMaxFperiod + kMaxLpf];
  __shared__ double s_lpf[2][kMaxLpf];
  __shared__ double s_c0[kMaxFperiod];
  const int nl = p.nlpf, fp = p.fperiod, hist = nl > 0 ? nl - 1 : 0;
  const int64_t n0 = (int64_t)f * fp;  // first sample of this frame within the utterance
  const bool voiced_cur = d.pitch[gf] != 0.0;
  const bool voiced_prev = f > 0 && d.pitch[gf - 1] != 0.0;

  for (int i = threadIdx.x; i < fp + hist; i += blockDim.x) {
    const int64_t m = n0 - hist + i;
    s_pulse[i] = m >= 0 ? d.pulse[foff * fp + m] : 0.0;
    s_noise[i] = m >= 0 ? d.noise[m] : 0.0;
  }
  for (int i = threadIdx.x; i < nl; i += blockDim.x) {
    s_lpf[1][i] = d.lpf[gf * nl + i];
    s_lpf[0][i] = f > 0 ? d.lpf[(gf - 1) * nl + i] : 0.0;
  }
  if (threadIdx.x == 0) {
    // coefficients[0] by repeated addition, as the per-sample loop does (mod.rs:133-135)
    const double cc = d.b[gf * p.nmcp];
    double c = f > 0 ? d.b[(gf - 1) * p.nmcp] : cc;
    const double cinc = (cc - c) / (double)fp;
    for (int i = 0; i < fp; ++i) {
      s_c0[i] = c;
      c += cinc;
    }
  }
  __syncthreads();

  for (int i = threadIdx.x; i < fp; i += blockDim.x) {
    double x;
    if (nl > 0) {
      const int h = (nl - 1) / 2;
      x = 0.0;
      for (int k = nl - 1; k >= 0; --k) {  // source m = n - k, oldest first
        const int j = i + hist - k;        // index into the staged window
        const int64_t m = n0 + i - k;
        if (m < 0) continue;
        const bool cur = (i - k) >= 0;     // source sample lies in this frame
        if (k == h) x += s_noise[j];
        if (cur ? voiced_cur : voiced_prev) x += (s_pulse[j] - s_noise[j]) * s_lpf[cur ? 1 : 0][k];
      }
    } else {
      x = voiced_cur ? s_pulse[i] : s_noise[i];
    }
    if (x != 0.0) x *= exp(s_c0[i]);
    d.wave[foff * fp + n0 + i] = x;
  }
}

// ---------------------------------------------------------------- MLSA, variant 0
// Bit-faithful to the FMA flavour of the reference (mlsa.rs:48-163). G lanes per utterance:
// lane i-1 (i = 1..5) owns Pade stage i: its df1 section (d11[i], d12[i]) and its 33-tap warped
// FIR delay line d21[i-1] in registers; the stage outputs are combined through warp shuffles in
// the reference's order (i = 5..1). Coefficients live in shared memory and advance by cinc once
// per sample.
template <int NM>  // NM = nmcp
__device__ __forceinline__ double fir_regs(double (&dl)[NM - 1], double x, double a, double aa,
                                           double aaaa, double iaa, const double* __restrict__ c,
                                           int cstride) {
  // dl[k-1] = d[k], k = 1..NM-1 ; coefficient k at c[k*cstride]
  double rem = -a * x + dl[0];
  dl[0] = iaa * x + a * dl[0];
  double y0 = 0.0, y1 = 0.0;
  constexpr int n = NM - 2, nchunk = n / 4;
#pragma unroll
  for (int q = 0; q < nchunk; ++q) {
    const int k = 2 + 4 * q;  // tap index of d0
    const double d0 = dl[k - 1], d1 = dl[k], d2 = dl[k + 1], d3 = dl[k + 2];
    const double t01 = __fma_rn(-a, d0, d1), t23 = __fma_rn(-a, d2, d3);
    const double tr0 = __fma_rn(-a, rem, d0), t12 = __fma_rn(-a, d1, d2);
    const double n0 = __fma_rn(iaa, rem, a * d0);
    const double n1 = __fma_rn(iaa, tr0, a * d1);
    const double n2 = __fma_rn(iaa, __fma_rn(aa, rem, t01), a * d2);
    const double n3 = __fma_rn(iaa, __fma_rn(aa, tr0, t12), a * d3);
    rem = __fma_rn(aaaa, rem, __fma_rn(aa, t01, t23));
    dl[k - 1] = n0;
    dl[k] = n1;
    dl[k + 1] = n2;
    dl[k + 2] = n3;
    y0 += c[(k)*cstride] * n0 + c[(k + 2) * cstride] * n2;
    y1 += c[(k + 1) * cstride] * n1 + c[(k + 3) * cstride] * n3;
  }
#pragma unroll
  for (int k = 2 + 4 * nchunk; k < NM; ++k) {
    const double dv = dl[k - 1];
    const double nd = iaa * rem + a * dv;
    rem = -a * rem + dv;
    dl[k - 1] = nd;
    y0 += c[k * cstride] * nd;
  }
  return y0 + y1;
}

template <int G, int NM, int WPB>
__global__ void __launch_bounds__(32 * WPB) k_mlsa_v0(BatchLayout b, VocoderParams p, VocoderDev d) {
  constexpr int UPW = 32 / G;        // utterances per warp
  constexpr int GPB = UPW * WPB;     // groups per block
  constexpr int CH = G < 16 ? G : 16;  // samples staged per chunk
  __shared__ double s_c[NM][GPB], s_cinc[NM][GPB], s_cc[NM][GPB];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int grp = lane / G, gl = lane % G;
  const int sg = warp * UPW + grp;  // group slot in shared memory
  const int u = (blockIdx.x * WPB + warp) * UPW + grp;
  const bool live = u < b.n_utt;
  const int uu = live ? u : 0;
  const int64_t foff = b.frame_off[uu];
  const int T = live ? (int)(b.frame_off[uu + 1] - foff) : 0;
  int f0 = d.frame_begin, f1 = f0 + d.frame_count;
  if (f1 > T || f1 < f0) f1 = T;
  if (f0 > T) f0 = T;
  // frames this warp must iterate = max over its groups (uniform control flow for shuffles)
  int nfr = f1 - f0;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) nfr = max(nfr, __shfl_xor_sync(0xffffffffu, nfr, o));

  const int st = gl + 1;            // Pade stage / section index owned by this lane (1..5 valid)
  const bool stage_lane = gl < 5;
  const double a = p.alpha, aa = a * a, aaaa = aa * aa, iaa = 1.0 - aa;
  const double pp = stage_lane ? c_ppade[st] : 0.0;
  const int base_lane = grp * G;

  double dl[NM - 1];
#pragma unroll
  for (int k = 0; k < NM - 1; ++k) dl[k] = 0.0;
  double d11 = 0.0, d12 = 0.0, d22 = 0.0;  // own section state
  double in1 = 0.0, in2 = 0.0;             // d12[st-1], d22[st-1] of the previous sample

  double* stp = d.state ? d.state + (int64_t)uu * kVocStateDoubles : nullptr;
  const bool resume = stp != nullptr && d.frame_begin > 0;
  if (resume && live) {
    if (stage_lane) {
#pragma unroll
      for (int k = 0; k < NM - 1; ++k) dl[k] = stp[gl * (NM - 1) + k];
      d11 = stp[170 + st];
      d12 = stp[176 + st];
      d22 = stp[182 + st];
      in1 = stp[176 + st - 1];
      in2 = stp[182 + st - 1];
    }
    for (int k = gl; k < NM; k += G) s_c[k][sg] = stp[188 + k];
  }
  __syncwarp();

  for (int fi = 0; fi < nfr; ++fi) {
    const int f = f0 + fi;
    const bool factive = live && f < f1;
    const int64_t gf = foff + (factive ? f : 0);
    // frame prologue (mod.rs:116-125): cc = mc2b(spectrum); cinc = (cc - c)/fperiod
    for (int k = gl; k < NM; k += G) {
      const double cc = factive ? d.b[gf * NM + k] : 0.0;
      const double c = (f == 0 || !factive) ? cc : s_c[k][sg];  // is_first: c = mc2b(first frame)
      s_c[k][sg] = c;
      s_cc[k][sg] = cc;
      s_cinc[k][sg] = (cc - c) / (double)p.fperiod;
    }
    __syncwarp();
    double* wv = d.wave + (foff + (factive ? f : 0)) * p.fperiod;
    for (int i0 = 0; i0 < p.fperiod; i0 += CH) {
      const int nch = min(CH, p.fperiod - i0);
      double xin = (factive && gl < nch) ? wv[i0 + gl] : 0.0;
      double xout = 0.0;
      for (int j = 0; j < nch; ++j) {
        double x = __shfl_sync(0xffffffffu, xin, base_lane + j);
        // ---- df1 (mlsa.rs:54-66)
        const double c1 = s_c[1][sg];
        d11 = iaa * in1 + a * d11;
        d12 = d11 * c1;
        double v = d12 * pp;
        double out = 0.0;
#pragma unroll
        for (int i = 5; i >= 1; --i) {
          const double vi = __shfl_sync(0xffffffffu, v, base_lane + i - 1);
          x += (i & 1) ? vi : -vi;
          out += vi;
        }
        {
          const double up = __shfl_up_sync(0xffffffffu, d12, 1);
          in1 = (gl == 0) ? x : up;  // d12[0] = x ; d12[st-1] for the next sample
        }
        x += out;
        // ---- df2 (mlsa.rs:69-79)
        d22 = fir_regs<NM>(dl, in2, a, aa, aaaa, iaa, &s_c[0][sg], GPB);
        v = d22 * pp;
        out = 0.0;
#pragma unroll
        for (int i = 5; i >= 1; --i) {
          const double vi = __shfl_sync(0xffffffffu, v, base_lane + i - 1);
          x += (i & 1) ? vi : -vi;
          out += vi;
        }
        {
          const double up = __shfl_up_sync(0xffffffffu, d22, 1);
          in2 = (gl == 0) ? x : up;
        }
        x += out;
        __syncwarp();
        for (int k = gl; k < NM; k += G) s_c[k][sg] += s_cinc[k][sg];  // mod.rs:133-135
        __syncwarp();
        if (gl == j) xout = x * p.volume;
      }
      if (factive && gl < nch) wv[i0 + gl] = xout;
    }
    for (int k = gl; k < NM; k += G) s_c[k][sg] = s_cc[k][sg];  // *coefficients = cc (mod.rs:140)
    __syncwarp();
  }

  if (stp && live) {
    if (stage_lane) {
#pragma unroll
      for (int k = 0; k < NM - 1; ++k) stp[gl * (NM - 1) + k] = dl[k];
      stp[170 + st] = d11;
      stp[176 + st] = d12;
      stp[182 + st] = d22;
      if (gl == 0) {
        stp[176] = in1;  // d12[0]
        stp[182] = in2;  // d22[0]
      }
    }
    for (int k = gl; k < NM; k += G) stp[188 + k] = s_c[k][sg];
  }
}

__global__ void k_gather_chunk(BatchLayout b, VocoderParams p, VocoderDev d, double* chunk) {
  const int u = blockIdx.y;
  const int64_t foff = b.frame_off[u];
  const int64_t T = b.frame_off[u + 1] - foff;
  const int64_t n = (int64_t)d.frame_count * p.fperiod;
  const int64_t avail = (T - d.frame_begin) * p.fperiod;  // may be <= 0
  const double* src = d.wave + (foff + d.frame_begin) * p.fperiod;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    chunk[(int64_t)u * n + i] = i < avail ? src[i] : 0.0;
}

}  // namespace

int launch_gather_chunk(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d,
                        double* chunk, cudaStream_t stream) {
  if (b.n_utt == 0 || d.frame_count <= 0) return 0;
  const int64_t n = (int64_t)d.frame_count * p.fperiod;
  dim3 grid((unsigned)std::min<int64_t>((n + 255) / 256, 64), (unsigned)b.n_utt);
  k_gather_chunk<<<grid, 256, 0, stream>>>(b, p, d, chunk);
  return (int)cudaGetLastError();
}

int launch_frame_prep(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d,
                      int64_t total_frames, cudaStream_t stream) {
  if (total_frames == 0) return 0;
  const int threads = 128;
  const int blocks = (int)((total_frames + threads - 1) / threads);
  k_frame_prep<<<blocks, threads, 0, stream>>>(p, d, total_frames);
  return (int)cudaGetLastError();
}

int launch_pulse(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, cudaStream_t stream) {
  if (b.n_utt == 0) return 0;
  const int threads = 32;
  k_pulse<<<(b.n_utt + threads - 1) / threads, threads, 0, stream>>>(b, p, d);
  return (int)cudaGetLastError();
}

int launch_excitation(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d,
                      int64_t total_frames, cudaStream_t stream) {
  if (total_frames == 0) return 0;
  if (p.nlpf > kMaxLpf || p.fperiod > kMaxFperiod) return (int)cudaErrorInvalidValue;
  k_excitation<<<(unsigned)total_frames, 256, 0, stream>>>(b, p, d, total_frames);
  return (int)cudaGetLastError();
}

int launch_mlsa(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, int variant,
                cudaStream_t stream) {
  if (b.n_utt == 0) return 0;
  if (p.nmcp != 35) return (int)cudaErrorInvalidValue;  // delay lines are register arrays
  (void)variant;
  constexpr int G = 8, WPB = 1;
  const int upb = (32 / G) * WPB;
  k_mlsa_v0<G, 35, WPB><<<(b.n_utt + upb - 1) / upb, 32 * WPB, 0, stream>>>(b, p, d);
  return (int)cudaGetLastError();
}

}  // namespace jbk
