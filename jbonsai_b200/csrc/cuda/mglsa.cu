// Stage != 0 arm of the vocoder (voices with the GAMMA option): line spectral pairs in, MGLSA filter.
//
// Reference: Vocoder::synthesize, Stage::NonZero arm (src/vocoder/mod.rs:89-107,142-176):
//   per frame   lsp -> postfilter_lsp(beta) -> check_lsp_stability -> lsp2mgc -> mc2b -> gnorm,
//               coefficients[1..] *= gamma                      (lsp.rs:27-165, cepstrum.rs:68-101,
//                                                                generalized.rs:6-38, gamma = -1/stage)
//   per sample  x = excitation * coefficients[0]; `stage` cascaded sections dff (mglsa.rs:15-41);
//               coefficients += cinc; out = x * volume
//
//   k_lsp_prep  one thread per frame (plus one per utterance for the un-filtered first frame the
//               reference starts from, mod.rs:92-107): the O(M^2) recurrences run in thread-local
//               arrays in the reference's operation order. cos/exp/ln/pow are CUDA's (<= 2 ulp from
//               glibc's), so the rows agree with the oracle to ~1e-13 relative, not bit for bit.
//   k_mglsa     the filter, time-segmented like MLSA variant 3 (mlsa_seg.cu): an utterance is cut
//               into segments that start `warm` frames early from a zero state; both sides of every
//               cut are stored and compared (k_mlsa_seg_check), and the host re-runs the kernel on
//               whole utterances when a cut disagrees. lane = STAGE * slot + section: the sections
//               of one segment form a pipeline one sample apart (section s filters sample n - s at
//               step n and hands its output to the next lane by shuffle), every lane keeps the
//               delay line d[0..M) of its section and its own copy of the interpolated coefficients
//               in registers (their per-sample increments sit in shared memory). Two flavours of the
//               tap update: LITERAL follows mglsa.rs:30-34 operation for operation (separate multiplies
//               and adds; with the oracle's coefficient rows and no warm-up it reproduces the oracle bit
//               for bit) and costs three dependent FP64 operations per tap; the default computes the same
//               quantity as (d[i] + alpha d[i+1]) - alpha d'[i-1] with fused multiply-adds, one operation
//               per tap on the serial path (the filter is a feedback loop: the chain over the taps is
//               what bounds a warp). The two differ by rounding only, ~1e-13 relative on the output.
#include "kernels.h"

namespace jbk {

namespace {

constexpr int kMaxM = 64;   // k_lsp_prep: largest vector length
constexpr int kMG = kMglsaMaxOrder;  // k_mglsa: largest vector length held in registers (36)
constexpr int kCH = 16;     // samples per staged chunk
constexpr int kWPB = 4;

__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gmem_src) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(sa), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// Generalized::gnorm / ignorm in place (generalized.rs:6-38), gamma != 0
__device__ void gnorm_dev(double* c, int n, double gamma) {
  const double k = 1.0 + gamma * c[0];
  c[0] = pow(k, 1.0 / gamma);
  for (int i = 1; i < n; ++i) c[i] = c[i] / k;
}
__device__ void ignorm_dev(double* c, int n, double gamma) {
  const double k = pow(c[0], gamma);
  c[0] = (k - 1.0) / gamma;
  for (int i = 1; i < n; ++i) c[i] = c[i] * k;
}

// LineSpectralPairs::lsp2mgc (lsp.rs:27-102): lsp[m] -> out[m]. The whole vector, gain slot included,
// is read as m line spectral frequencies, exactly as the reference does.
__device__ void lsp2mgc_dev(const double* lsp, int m, bool use_log_gain, int stage, double gamma, double* out) {
  double a0[kMaxM / 2 + 2], a1[kMaxM / 2 + 2], a2[kMaxM / 2 + 2];
  double b0[kMaxM / 2 + 2], b1[kMaxM / 2 + 2], b2[kMaxM / 2 + 2];
  double pc[kMaxM / 2 + 1], qc[kMaxM / 2 + 1];
  double lpc[kMaxM + 1];
  const bool odd = (m & 1) != 0;
  const int mh1 = odd ? (m + 1) / 2 : m / 2, mh2 = odd ? (m - 1) / 2 : m / 2;
  for (int i = 0; i < mh1; ++i) pc[i] = -2.0 * cos(lsp[2 * i]);
  for (int i = 0; i < mh2; ++i) qc[i] = -2.0 * cos(lsp[2 * i + 1]);
  for (int i = 0; i <= mh1; ++i) a0[i] = a1[i] = a2[i] = 0.0;
  for (int i = 0; i <= mh2; ++i) b0[i] = b1[i] = b2[i] = 0.0;
  double xff = 0.0, xf = 0.0;
  for (int k = 0; k <= m; ++k) {
    const double xx = k == 0 ? 1.0 : 0.0;
    if (odd) {
      a0[0] = xx;
      b0[0] = xx - xff;
      xff = xf;
      xf = xx;
    } else {
      a0[0] = xx + xf;
      b0[0] = xx - xf;
      xf = xx;
    }
    for (int i = 0; i < mh1; ++i) {
      a0[i + 1] = a0[i] + pc[i] * a1[i] + a2[i];
      a2[i] = a1[i];
      a1[i] = a0[i];
    }
    for (int i = 0; i < mh2; ++i) {
      b0[i + 1] = b0[i] + qc[i] * b1[i] + b2[i];
      b2[i] = b1[i];
      b1[i] = b0[i];
    }
    if (k > 0) lpc[k - 1] = -0.5 * (a0[mh1] + b0[mh2]);
  }
  for (int i = m - 1; i >= 0; --i) lpc[i + 1] = -lpc[i];
  // lsp.rs:92-100
  lpc[0] = use_log_gain ? exp(lsp[0]) : lsp[0];
  ignorm_dev(lpc, m + 1, gamma);
  for (int i = 1; i <= m; ++i) lpc[i] *= -(double)stage;
  // mgc2mgc with equal alphas (cepstrum.rs:94-96): gnorm, gc2gc(m - 1, gamma), ignorm
  gnorm_dev(lpc, m + 1, gamma);
  const int m2 = m - 1, n1 = m + 1;
  out[0] = lpc[0];
  for (int i = 1; i <= m2; ++i) {
    double ss1 = 0.0, ss2 = 0.0;
    const int kmax = n1 < i ? n1 : i;
    for (int k = 1; k < kmax; ++k) {
      const int mk = i - k;
      const double cc = lpc[k] * out[mk];
      ss1 += (double)mk * cc;
      ss2 += (double)k * cc;
    }
    const double corr = (gamma * ss2 - gamma * ss1) / (double)i;
    out[i] = i < n1 ? lpc[i] + corr : corr;
  }
  ignorm_dev(out, m, gamma);
}

__device__ double lsp2en_dev(const double* lsp, int m, bool use_log_gain, int stage, double gamma) {
  double mgc[kMaxM];
  lsp2mgc_dev(lsp, m, use_log_gain, stage, gamma, mgc);
  double e = 0.0;
  for (int i = 0; i < m; ++i) e += mgc[i] * mgc[i];
  return e;
}

// gf < total_frames: row gf of d.b from frame gf (post-filter and stability check applied);
// gf >= total_frames: row (gf - total_frames) of b_first from the first frame of that utterance as it is.
__global__ void __launch_bounds__(64)
k_lsp_prep(BatchLayout b, VocoderParams p, VocoderDev d, double beta, double* b_first, int64_t total_frames) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total_frames + b.n_utt) return;
  const int m = p.nmcp, stage = p.stage;
  const bool lg = p.use_log_gain != 0;
  const double gamma = -1.0 / (double)stage;
  const bool filtered = gid < total_frames;
  int64_t src = gid;
  double* dst;
  if (filtered) {
    dst = d.b + gid * m;
  } else {
    const int u = (int)(gid - total_frames);
    if (b.frame_off[u + 1] <= b.frame_off[u]) return;
    src = b.frame_off[u];
    dst = b_first + (int64_t)u * m;
  }
  double lsp[kMaxM];
  for (int i = 0; i < m; ++i) lsp[i] = d.mcp[src * m + i];

  if (filtered) {
    if (beta > 0.0 && m > 2) {  // postfilter_lsp (lsp.rs:108-137)
      double buf[kMaxM];
      const double en1 = lsp2en_dev(lsp, m, lg, stage, gamma);
      for (int i = 0; i < m; ++i) {
        if (i > 1 && i < m - 1) {
          const double d1 = beta * (lsp[i + 1] - lsp[i]);
          const double d2 = beta * (lsp[i] - lsp[i - 1]);
          buf[i] = lsp[i - 1] + d2 + (d2 * d2 * ((lsp[i + 1] - lsp[i - 1]) - (d1 + d2))) / ((d2 * d2) + (d1 * d1));
        } else {
          buf[i] = lsp[i];
        }
      }
      for (int i = 0; i < m; ++i) lsp[i] = buf[i];
      const double en2 = lsp2en_dev(lsp, m, lg, stage, gamma);
      if (en1 != en2) {
        if (lg) lsp[0] += 0.5 * log(en1 / en2);
        else lsp[0] *= sqrt(en1 / en2);
      }
    }
    {  // check_lsp_stability (lsp.rs:139-165)
      const double pi = 3.14159265358979323846;
      const double mn = 0.25 * pi / (double)m;
      const int last = m - 1;
      for (int it = 0; it < 4; ++it) {
        bool find = false;
        for (int j = 1; j < last; ++j) {
          const double tmp = lsp[j + 1] - lsp[j];
          if (tmp < mn) {
            lsp[j] -= 0.5 * (mn - tmp);
            lsp[j + 1] += 0.5 * (mn - tmp);
            find = true;
          }
        }
        if (lsp[1] < mn) {
          lsp[1] = mn;
          find = true;
        }
        if (lsp[last] > pi - mn) {
          lsp[last] = pi - mn;
          find = true;
        }
        if (!find) break;
      }
    }
  }
  double mgc[kMaxM];
  lsp2mgc_dev(lsp, m, lg, stage, gamma, mgc);
  // mc2b (cepstrum.rs:139-149), gnorm, coefficients[1..] *= gamma (mod.rs:151-154)
  if (p.alpha != 0.0)
    for (int i = m - 2; i >= 0; --i) mgc[i] = mgc[i] - p.alpha * mgc[i + 1];
  gnorm_dev(mgc, m, gamma);
  dst[0] = mgc[0];
  for (int i = 1; i < m; ++i) dst[i] = mgc[i] * gamma;
}

// binc[f][k] = (b[f][k] - b[f-1][k]) / fperiod, the per-sample increment of vocoder/mod.rs:156-160; frame 0 of
// an utterance starts from its row of b_first. One thread per element.
__global__ void k_row_increments(BatchLayout b, VocoderParams p, VocoderDev d, int64_t total_frames) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int M = p.nmcp;
  if (i >= total_frames * M) return;
  const int64_t gf = i / M;
  const int k = (int)(i - gf * M);
  int lo = 0, hi = b.n_utt;  // largest u with frame_off[u] <= gf
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (b.frame_off[mid] <= gf) lo = mid; else hi = mid;
  }
  // without b_first (stage 0, beta == 0) the first frame starts from its own row: no increment
  const double prev = gf == b.frame_off[lo] ? (d.b_first ? d.b_first[(int64_t)lo * M + k] : d.b[i]) : d.b[i - M];
  d.binc[i] = (d.b[i] - prev) / (double)p.fperiod;
}

// ------------------------------------------------------------------------------------------ filter
struct SlotInfo {
  long long in_base;   // index of the first sample this segment filters (frame `warm`)
  long long b_base;    // frame_off[utt] * nmcp
  int warm, last_frame, nfr, own0, seg, prev, utt;
};

template <int STAGE>
struct WarpShared {
  static constexpr int kSlots = 32 / STAGE;
  double In[2 * kCH * kSlots];   // excitation ring: [sample mod 32][slot]
  double Out[2 * kCH * kSlots];  // output ring, the last section runs STAGE - 1 samples behind the first
  double cinc[(kMG - 1) * 32];   // per-sample coefficient increments, [k][lane]
  SlotInfo info[kSlots];
};

// MT > 0: the vector length is the compile-time constant MT (the tap loop is straight-line code);
// MT == 0: any length up to kMG, decided at run time (a branch per tap).
template <int STAGE, bool LITERAL, int MT>
__global__ void __launch_bounds__(32 * kWPB, 2)
k_mglsa(BatchLayout b, VocoderParams p, VocoderDev d, MlsaSegments sg, int seg_first, int seg_end, int state_stride,
        const double* __restrict__ binc) {
  constexpr int kSlots = 32 / STAGE;
  extern __shared__ __align__(16) unsigned char mg_smem[];
  WarpShared<STAGE>& S = reinterpret_cast<WarpShared<STAGE>*>(mg_smem)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const int warp = blockIdx.x * kWPB + (threadIdx.x >> 5);
  if (seg_first + warp * kSlots >= seg_end) return;  // whole warp; warps of a block never sync with each other
  SlotInfo* const info = S.info;
  const int fp = p.fperiod, M = MT > 0 ? MT : p.nmcp;

  for (int k = lane; k < kSlots; k += 32) {
    SlotInfo si;
    const int sid = seg_first + warp * kSlots + k;
    if (sid < seg_end) {
      const MlsaSegment s = sg.seg[sid];
      const long long foff = b.frame_off[s.utt];
      si.in_base = (foff + s.warm) * (long long)fp;
      si.b_base = foff * M;
      si.warm = s.warm;
      si.last_frame = (int)(b.frame_off[s.utt + 1] - foff) - 1;
      si.nfr = s.end - s.warm;
      si.own0 = s.begin - s.warm;
      si.seg = sid;
      si.prev = s.prev;
      si.utt = s.utt;
    } else {
      si.in_base = 0; si.b_base = 0; si.warm = 0; si.last_frame = 0; si.nfr = 0; si.own0 = 0; si.seg = -1;
      si.prev = -1; si.utt = 0;
    }
    info[k] = si;
  }
  __syncwarp();

  const bool live = lane < kSlots * STAGE;  // 32 % STAGE lanes at the top have no section to run
  const int g = live ? lane / STAGE : kSlots - 1, s = live ? lane - g * STAGE : 0;
  const bool tail = live && s == STAGE - 1;
  const double a = p.alpha, ma = -a, aa = 1.0 - a * a;
  const SlotInfo me = info[g];
  int nfr_max = me.nfr;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) nfr_max = max(nfr_max, __shfl_xor_sync(0xffffffffu, nfr_max, o));
  if (nfr_max <= 0) return;

  // c[k] = coefficients[k + 1]; coefficients[0] is the gain k_excitation has already applied
  const double* const brow = d.b + me.b_base + 1;
  auto frame_row = [&](int frame_rel) -> const double* {
    int f = me.warm + frame_rel;
    if (f < 0) return d.b_first + (long long)me.utt * M + 1;  // the un-filtered first frame (mod.rs:92-107)
    f = f > me.last_frame ? me.last_frame : f;
    return brow + (long long)f * M;
  };
  double c[kMG - 1], dl[kMG];
  double* const cinc = S.cinc + lane;  // [k * 32]: registers hold the delay line and the coefficients only
#pragma unroll
  for (int k = 0; k < kMG - 1; ++k) { c[k] = 0.0; cinc[k * 32] = 0.0; }
#pragma unroll
  for (int k = 0; k < kMG; ++k) dl[k] = 0.0;
  int next_frame = 0;  // frame (relative to warm) whose row the coefficients move towards next
  auto frame_boundary = [&]() {  // mod.rs:156-160 with coefficients = cc of the frame before (:175)
    const double* r0 = frame_row(next_frame - 1);
    // (cc - coefficients) / fperiod of every frame comes from k_row_increments; past the end of the
    // utterance (a shorter segment idling beside a longer one) the coefficients stand still
    const int f = me.warm + next_frame;
    const double* ri = binc + me.b_base + 1 + (long long)(f > me.last_frame ? me.last_frame : f) * M;
    const bool past = f > me.last_frame;
    ++next_frame;
#pragma unroll
    for (int k = 0; k < kMG - 1; ++k)
      if (k < M - 1) {
        c[k] = r0[k];
        cinc[k * 32] = past ? 0.0 : ri[k];
      }
  };
  // section s starts s steps late: until then it filters zeros and its coefficients stand still
  int left = s == 0 ? fp : s;
  if (s == 0) frame_boundary();

  const int nsmp_max = nfr_max * fp;
  const int nsteps = nsmp_max + STAGE - 1;
  auto fill_chunk = [&](int chunk) {
    for (int k = lane; k < kSlots * kCH; k += 32) {
      const int gg = k / kCH, i = k - gg * kCH;
      const SlotInfo& si = info[gg];
      const int n = chunk * kCH + i;
      double* dst = S.In + ((chunk & 1) * kCH + i) * kSlots + gg;
      if (n < si.nfr * fp) cp_async8(dst, d.wave + si.in_base + n);  // lands while the next chunk is filtered
      else *dst = 0.0;
    }
  };
  fill_chunk(0);
  fill_chunk(1);
  cp_async_wait_all();
  __syncwarp();

  const bool my_cut = me.prev >= 0 && live && me.seg >= 0 && sg.state != nullptr;
  const bool my_end = live && me.seg >= 0 && sg.state != nullptr;
  constexpr int kNever = -(1 << 30);  // ks starts at -s: a negative sample index must not match
  const int dump0 = my_cut ? me.own0 * fp : kNever, dump1 = my_end ? me.nfr * fp : kNever;
  auto dump_state = [&](int which) {
    double* o = sg.state + ((long long)me.seg * 2 + which) * state_stride + s * kMG;
#pragma unroll
    for (int i = 0; i < kMG; ++i) o[i] = dl[i];
  };

  int ks = -s;  // index of the sample this lane filters at the current step
  double xout = 0.0;
  for (int i0 = 0; i0 < nsteps; i0 += kCH) {
#pragma unroll 1
    for (int j = 0; j < kCH; ++j) {
      if (ks == dump0) dump_state(0);
      if (ks == dump1) dump_state(1);
      const double xprev = __shfl_up_sync(0xffffffffu, xout, 1);
      double x = s == 0 ? S.In[((i0 + j) & (2 * kCH - 1)) * kSlots + g] : xprev;
      // dff (mglsa.rs:21-41); the shift d[i] = d[i-1] is folded into the walk
      const double d0 = dl[0];
      double y = d0 * c[0];
      double prev_new = d0, old_next = dl[1];
#pragma unroll
      for (int i = 1; i < kMG; ++i) {
        if (i <= M - 2) {
          const double cur_old = old_next;
          old_next = dl[i + 1 < kMG ? i + 1 : i];
          double nw;
          if (LITERAL) {  // d[i] += alpha * (d[i+1] - d[i-1]); y += d[i] * c[i+1]   (mglsa.rs:31-34)
            nw = __dadd_rn(cur_old, __dmul_rn(a, __dadd_rn(old_next, -prev_new)));
            y = __dadd_rn(y, __dmul_rn(nw, c[i]));
          } else {  // the same value as (d[i] + alpha d[i+1]) - alpha d'[i-1]: one fused operation on the serial path
            nw = __fma_rn(ma, prev_new, __fma_rn(a, old_next, cur_old));
            y = __fma_rn(nw, c[i], y);
          }
          c[i] = __dadd_rn(c[i], cinc[i * 32]);  // coefficient of the next sample (mod.rs:169-171)
          dl[i] = prev_new;
          prev_new = nw;
        } else if (i == M - 1) {
          dl[i] = prev_new;
        }
      }
      x = __dadd_rn(x, -y);
      dl[0] = __dadd_rn(__dmul_rn(a, d0), __dmul_rn(aa, x));
      xout = x;
      if (tail && ks >= 0) S.Out[(ks & (2 * kCH - 1)) * kSlots + g] = x;
      c[0] = __dadd_rn(c[0], cinc[0]);
      if (--left == 0) {  // a frame end replaces the coefficients (mod.rs:175)
        left = fp;
        frame_boundary();
      }
      ++ks;
    }
    cp_async_wait_all();  // chunk i0 / kCH + 1 has landed
    __syncwarp();
    {  // flush the samples whose last section has run; refill the half of the ring just consumed
      const int chunk = i0 / kCH;
      for (int k = lane; k < kSlots * kCH; k += 32) {
        const int gg = k / kCH, i = k - gg * kCH;
        const SlotInfo& si = info[gg];
        const int n = i0 - (STAGE - 1) + i;
        if (n >= 0 && n >= si.own0 * fp && n < si.nfr * fp)
          d.out[si.in_base + n] = S.Out[(n & (2 * kCH - 1)) * kSlots + gg] * p.volume;
      }
      if ((chunk + 2) * kCH < nsteps) fill_chunk(chunk + 2);
    }
    __syncwarp();
  }
  if (ks == dump1) dump_state(1);  // one section: the last step of the loop is the last sample
}

template <int STAGE, bool LITERAL, int MT>
int launch_one(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, const MlsaSegments& sg,
               int seg_first, int seg_count, cudaStream_t stream) {
  constexpr int kSlots = 32 / STAGE;
  const int warps = (seg_count + kSlots - 1) / kSlots;
  const size_t smem = sizeof(WarpShared<STAGE>) * kWPB;
  cudaError_t e =
      cudaFuncSetAttribute(k_mglsa<STAGE, LITERAL, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  k_mglsa<STAGE, LITERAL, MT><<<(warps + kWPB - 1) / kWPB, 32 * kWPB, smem, stream>>>(
      b, p, d, sg, seg_first, seg_first + seg_count, mglsa_state_doubles(p.stage), d.binc);
  return (int)cudaGetLastError();
}
template <int STAGE>
int launch_either(bool literal, const BatchLayout& b, const VocoderParams& p, const VocoderDev& d,
                  const MlsaSegments& sg, int seg_first, int seg_count, cudaStream_t stream) {
  if (literal) return launch_one<STAGE, true, 0>(b, p, d, sg, seg_first, seg_count, stream);
  if (p.nmcp == 35) return launch_one<STAGE, false, 35>(b, p, d, sg, seg_first, seg_count, stream);  // the shipped order
  return launch_one<STAGE, false, 0>(b, p, d, sg, seg_first, seg_count, stream);
}

}  // namespace

int launch_lsp_prep(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, double beta, double* b_first,
                    int64_t total_frames, cudaStream_t stream) {
  if (total_frames == 0) return 0;
  if (p.stage <= 0 || p.nmcp < 3 || p.nmcp > kMaxM || b_first == nullptr) return (int)cudaErrorInvalidValue;
  const int64_t n = total_frames + b.n_utt;
  k_lsp_prep<<<(unsigned)((n + 63) / 64), 64, 0, stream>>>(b, p, d, beta, b_first, total_frames);
  return (int)cudaGetLastError();
}

int launch_row_increments(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, int64_t total_frames,
                          cudaStream_t stream) {
  if (total_frames == 0) return 0;
  if (d.binc == nullptr) return (int)cudaErrorInvalidValue;
  const int64_t n = total_frames * p.nmcp;
  k_row_increments<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(b, p, d, total_frames);
  return (int)cudaGetLastError();
}

namespace {
// Sequential MGLSA with carried state: SpeechGenerator::generate_step (speech.rs:65-83) driving the Stage::NonZero arm
// (vocoder/mod.rs:142-176) a few frames at a time, and any frame period. One thread per utterance filters frames
// [frame_begin, frame_begin + frame_count) in place in d.wave: per sample the `stage` sections of mglsa.rs:15-19 one
// after the other in the reference's operation order (dff, mglsa.rs:21-41), then the coefficient step
// (mod.rs:169-171). The delay lines live in `state` ([n_utt][stage][kMG], local copy while the chunk runs); the
// coefficients need no state (they restart from the row of the frame before at every frame, mod.rs:175).
__global__ void __launch_bounds__(64) k_mglsa_stream(BatchLayout b, VocoderParams p, VocoderDev d, double* state) {
  int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= b.n_utt) return;
  if (b.order) u = b.order[u];
  const int64_t foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  int f0 = d.frame_begin, f1 = f0 + d.frame_count;
  if (f1 > T || f1 < f0) f1 = T;
  if (f0 >= f1) return;
  const int M = p.nmcp, fp = p.fperiod, NS = p.stage;
  const double a = p.alpha, aa = 1.0 - a * a;
  double dl[kMglsaMaxStage][kMG];
  double* const st = state ? state + (int64_t)u * kMglsaMaxStage * kMG : nullptr;
  for (int s = 0; s < NS; ++s)
    for (int i = 0; i < kMG; ++i) dl[s][i] = (st && f0 > 0) ? st[s * kMG + i] : 0.0;
  double c[kMG], cinc[kMG];
  for (int f = f0; f < f1; ++f) {
    const double* r0 = f == 0 ? d.b_first + (int64_t)u * M : d.b + (foff + f - 1) * M;  // mod.rs:92-107 / :175
    const double* ri = d.binc + (foff + f) * M;
    for (int k = 0; k < M; ++k) {
      c[k] = r0[k];
      cinc[k] = ri[k];
    }
    double* wv = d.wave + (foff + f) * fp;
    for (int n = 0; n < fp; ++n) {
      double x = wv[n];  // excitation times coefficients[0] (k_excitation, mod.rs:163)
      for (int s = 0; s < NS; ++s) {  // dff (mglsa.rs:21-41)
        double* dd = dl[s];
        const double d0 = dd[0];
        double y = __dmul_rn(d0, c[1]);
        double prev_new = d0, old_next = dd[1];
        for (int i = 1; i <= M - 2; ++i) {
          const double cur_old = old_next;
          old_next = dd[i + 1 < kMG ? i + 1 : i];
          const double nw = __dadd_rn(cur_old, __dmul_rn(a, __dadd_rn(old_next, -prev_new)));
          y = __dadd_rn(y, __dmul_rn(nw, c[i + 1]));
          dd[i] = prev_new;
          prev_new = nw;
        }
        dd[M - 1] = prev_new;
        x = __dadd_rn(x, -y);
        dd[0] = __dadd_rn(__dmul_rn(a, d0), __dmul_rn(aa, x));
      }
      for (int k = 0; k < M; ++k) c[k] = __dadd_rn(c[k], cinc[k]);
      wv[n] = x * p.volume;
    }
  }
  if (st)
    for (int s = 0; s < NS; ++s)
      for (int i = 0; i < kMG; ++i) st[s * kMG + i] = dl[s][i];
}
}  // namespace

int launch_mglsa_stream(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, double* state,
                        cudaStream_t stream) {
  if (b.n_utt == 0) return 0;
  if (p.nmcp < 3 || p.nmcp > kMG || p.stage < 1 || p.stage > kMglsaMaxStage || d.b_first == nullptr || d.binc == nullptr)
    return (int)cudaErrorInvalidValue;
  k_mglsa_stream<<<(b.n_utt + 63) / 64, 64, 0, stream>>>(b, p, d, state);
  return (int)cudaGetLastError();
}

int launch_mglsa(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, const MlsaSegments& sg,
                 int seg_first, int seg_count, bool literal, cudaStream_t stream) {
  if (seg_count <= 0) return 0;
  if (p.nmcp < 3 || p.nmcp > kMG || p.fperiod % kCH != 0 || d.out == nullptr || d.out == d.wave ||
      d.b_first == nullptr || d.binc == nullptr || seg_first < 0 || seg_first + seg_count > sg.n_seg)
    return (int)cudaErrorInvalidValue;
  switch (p.stage) {
    case 1: return launch_either<1>(literal, b, p, d, sg, seg_first, seg_count, stream);
    case 2: return launch_either<2>(literal, b, p, d, sg, seg_first, seg_count, stream);
    case 3: return launch_either<3>(literal, b, p, d, sg, seg_first, seg_count, stream);
    case 4: return launch_either<4>(literal, b, p, d, sg, seg_first, seg_count, stream);
    case 5: return launch_either<5>(literal, b, p, d, sg, seg_first, seg_count, stream);
    case 6: return launch_either<6>(literal, b, p, d, sg, seg_first, seg_count, stream);
    case 7: return launch_either<7>(literal, b, p, d, sg, seg_first, seg_count, stream);
    case 8: return launch_either<8>(literal, b, p, d, sg, seg_first, seg_count, stream);
    default: return (int)cudaErrorInvalidValue;
  }
}

}  // namespace jbk
