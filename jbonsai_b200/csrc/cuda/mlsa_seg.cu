// MLSA filter, variant 3 (default): time-segmented, six segments per warp, one lane per Pade stage.
//
// Reference: MelLogSpectrumApproximation::<6>::{df, df1, df2, fir} (src/vocoder/mlsa.rs:48-163)
// driven per sample by Vocoder::synthesize (src/vocoder/mod.rs:112-141).
//
// Why segments. The filter is a feedback loop along time, so one utterance is one serial chain
// and a batch of 1024 utterances can keep at most 1024 chains busy on a chip with 592 warp
// schedulers: variants 0-2 are latency-bound (profiles/r1_mlsa_v2_full_summary.txt). The filter's
// memory fades, though: started from a zero state W frames early, its state is indistinguishable
// from the true one (to the rounding-noise floor, ~1e-11 on +-3e4 samples) after about 32 frames
// of the shipped voices (tools/mlsa_memory_probe.py). The per-sample coefficients need no history
// at all (they restart from mc2b of the previous frame at every frame boundary, mod.rs:140). So an
// utterance is cut into segments; every segment but the first runs `warm` extra frames in front of
// its own range with its output discarded. Nothing is assumed: every segment stores its state at
// the start of its own range and at its end, k_mlsa_seg_check compares the two sides of every cut,
// and the host re-runs the sequential kernel (variant 2) when a cut disagrees by more than 1e-7
// (Engine::mlsa_cut_tolerance; the rounding-noise floor of that comparison is ~3e-8).
//
// Layout. lane = 5 g + s: segment slot g = 0..5 of this warp, Pade stage s + 1 (s = 0..4); lanes
// 30/31 shadow slot 5 and write to dump slots. A lane owns the whole 34-tap warped delay line of
// its stage in registers and runs the tap recurrence serially in the "q form"
//   t'_1 = (1-a^2) u + a t_1 ;  t'_j = (t_{j-1} + a t_j) - a t'_{j-1}   (== iaa*rem + a*d of mlsa.rs:141-153)
// with two dot-product accumulators like the reference (y0/y1). No scan, no carry exchange: three
// FMAs per tap, which is the arithmetic floor of the recurrence. Every lane also keeps the 34
// interpolated coefficients of its segment and their per-sample increments in registers and
// advances them itself (mod.rs:133-135): sharing them through shared memory cost more issue slots
// than the repeated additions. One shared-memory exchange per sample carries the five stage
// outputs and the five df1 outputs to the lanes of the segment; the Pade combine
// (d[0] = (x + A) - B, x' = x + 2A with A/B the odd/even weighted sums; mlsa.rs:54-79) is done
// redundantly by each lane.
//
// Measured limits (tools/ubench3.cu, B200): a DFMA whose three sources are all distinct registers
// issues every 3 cycles per scheduler, one with a reused/constant source every 2 (register-file
// read bandwidth). Per tap that is 2 + 2 + 3 (dot product) + 2 (coefficient add) = 9 issue cycles,
// ~350 per warp and sample. The kernel is bound by that issue rate, not by the latency of the tap
// chain: halving the chain with a two-tap look-ahead (one more FMA per pair) made it 12 % slower,
// exactly the growth in FP64 instructions.
#include <type_traits>

// 1: read the per-frame coefficient increments from VocoderDev::binc (launch_row_increments) instead of
// dividing at every frame boundary. Measured on B200: the kernel shrinks from ~2500 to ~900 instructions but
// runs 5 % slower (32.6 vs 31.0 ms per step; ptxas lands on 255 registers with a spill in the sample loop),
// so the divisions stay. The MGLSA kernel (mglsa.cu), with fewer live registers, does use the increments.
// 1: keep the per-sample coefficient increments in shared memory ([tap][lane], conflict-free) instead of 68
// registers, and build for three blocks per SM. Measured on B200: 240 registers and two warps per scheduler:
// MLSA 25.4 instead of 21.4 ms (34 more 64-bit shared loads per sample, two wavefronts each); 168 registers and
// three warps per scheduler (engine.cc kSegWarpsPerSmsp = 3): 116 B of spills, 37 % warm-up, 47 ms per step.
#ifndef JB_V3_CINC_SMEM
#define JB_V3_CINC_SMEM 0
#endif
#ifndef JB_V3_BINC
#define JB_V3_BINC 0
#endif
// 1 (default): the mc2b rows a frame boundary needs are staged in shared memory by cp.async one frame ahead.
// With 254 live registers ptxas cannot keep more than a few global loads in flight at the boundary, so the 68
// row loads ran one after another at L2/HBM latency: ~15 % of the kernel's stall samples sat in that block
// (ncu source page, r1e), for code that runs once per 240 samples.
#ifndef JB_V3_ROWS_SMEM
#define JB_V3_ROWS_SMEM 1
#endif
// 1: the first own0 - f64_tail warm-up frames of a warp run in FP32 (launch_mlsa_v3's f64_tail). Measured on
// B200 (bench batch, W = 40): the FP32 head leaves ~0.6 absolute in the worst cut's taps (states reach 1e4;
// rounding accumulates over the filter's memory) and the worst cut sheds only 0.19 decades per FP64 frame:
// mismatch 1.2e-5 / 3.6e-4 / 5.5e-3 / 0.11 with 20 / 12 / 8 / 4 double-precision frames behind it, against a
// tolerance of 1e-7 - it would take 36 of the 40 frames. Off.
#ifndef JB_V3_F32_HEAD
#define JB_V3_F32_HEAD 0
#endif
// How many of the 34 coefficient additions stay in front of the exchange loads (34: all, the first layout).
// The additions do not depend on the exchange, so the ones behind the loads fill the loads' latency; each
// load kept in flight across them costs two registers of a kernel that sits at the cap.
// 1: the frame-end block sits inside the sample step (first layout); 0: after the chunk's sample loop.
#ifndef JB_V3_BOUNDARY_IN_STEP
#define JB_V3_BOUNDARY_IN_STEP 1
#endif
#ifndef JB_V3_CADD_FRONT
#define JB_V3_CADD_FRONT 34
#endif

#include "kernels.h"

namespace jbk {

namespace {

__constant__ double c_pp3[6] = {1.00000000000, 0.49993910000, 0.11070980000,
                                0.01369984000, 0.00095648530, 0.00003041721};  // mlsa.rs:25-32

__device__ __forceinline__ void cp_async8s(double* smem_dst, const double* gmem_src) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(sa), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all_s() {
  asm volatile("cp.async.wait_all;\n" ::: "memory");
}

constexpr int kNM = 35;       // mel-cepstral order + 1
constexpr int kTaps = 34;     // delay-line taps d[1..34] (d[0] is write-only in the reference)
constexpr int kSlots = 6;     // segments per warp
constexpr int kCH = 16;       // samples per staged chunk (input ring of 2 chunks, output tile of 1)
constexpr int kWPB = 4;       // warps per block (independent of each other)

struct SlotInfo {
  long long in_base;   // index of the first sample this segment filters (frame `warm`)
  long long b_base;    // frame_off[utt] * kNM
  int warm;            // first filtered frame
  int last_frame;      // utterance length - 1
  int nfr;             // frames filtered = end - warm (0: empty slot)
  int own0;            // begin - warm: frames whose output is discarded
  int seg;             // segment id (state slot), -1 for an empty slot
  int prev;            // id of the segment before this one in the utterance, -1 if none
  int utt;             // utterance (row of VocoderDev::b_first)
};

// Shared memory of one warp. Every array is indexed so that the 16 lanes of a half-warp touch 16
// different 8-byte banks: Y/F by lane (= 5 slot + stage), In/Out as [sample][slot] rows of 8.
struct WarpShared {
  double Y[2][32];            // stage outputs y_1..y_5 of every slot, by lane; double-buffered by sample parity
  double F[2][32];            // df1 section outputs
  double In[2 * kCH * 8];     // excitation ring: [sample mod 32][slot]
  double Out[kCH * 8];        // output tile: [sample mod 16][slot]
#if JB_V3_ROWS_SMEM
  double R[2][kTaps * 8];     // mc2b rows of frames f and f + 1 of every slot, [frame parity][tap][slot]
#endif
  SlotInfo info[kSlots];
#if JB_V3_CINC_SMEM
  double cinc[kTaps * 32];    // per-sample coefficient increments of every lane, [tap][lane]
#endif
};

template <int WPB>
__global__ void __launch_bounds__(32 * WPB, JB_V3_CINC_SMEM ? 3 : 2)
k_mlsa_v3(BatchLayout b, VocoderParams p, VocoderDev d, MlsaSegments sg, int seg_first, int seg_end,
          const double iaa, const int f64_tail) {
#if JB_V3_CINC_SMEM
  extern __shared__ __align__(16) unsigned char v3_smem[];
  WarpShared* const sh = reinterpret_cast<WarpShared*>(v3_smem);
#else
  __shared__ __align__(16) WarpShared sh[WPB];
#endif

  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int warp = blockIdx.x * WPB + wib;
  if (seg_first + warp * kSlots >= seg_end) return;  // whole warp leaves; warps of a block never sync with each other
  WarpShared& S = sh[wib];
  SlotInfo* const info = S.info;

  const int fp = p.fperiod;
  // ---- slot table
  if (lane < kSlots) {
    SlotInfo si;
    const int sid = seg_first + warp * kSlots + lane;
    if (sid < seg_end) {
      const MlsaSegment s = sg.seg[sid];
      const long long foff = b.frame_off[s.utt];
      si.in_base = (foff + s.warm) * (long long)fp;
      si.b_base = foff * kNM;
      si.warm = s.warm;
      si.last_frame = (int)(b.frame_off[s.utt + 1] - foff) - 1;
      si.nfr = s.end - s.warm;
      si.own0 = s.begin - s.warm;
      si.seg = sid;
      si.prev = s.prev;
      si.utt = s.utt;
    } else {
      si.in_base = 0; si.b_base = 0; si.warm = 0; si.last_frame = 0; si.nfr = 0; si.own0 = 0; si.seg = -1; si.prev = -1;
      si.utt = 0;
    }
    info[lane] = si;
  }
  S.Y[0][lane] = 0.0; S.Y[1][lane] = 0.0; S.F[0][lane] = 0.0; S.F[1][lane] = 0.0;
  __syncwarp();

  const int g = lane < 30 ? lane / 5 : 5, s = lane < 30 ? lane - 5 * g : lane - 30;
  const int g5 = 5 * g;
  const int prev_lane = s > 0 ? g5 + s - 1 : g5;
  const bool head = lane < 30 && s == 0;
  const double a = p.alpha, ma = -a;  // iaa = 1 - a^2 is a kernel parameter (see the launcher)
  const double pp1 = c_pp3[1], pp2 = c_pp3[2], pp3 = c_pp3[3], pp4 = c_pp3[4], pp5 = c_pp3[5];
  const double fper = (double)fp;
  const SlotInfo me = info[g];
  const bool my_cut = me.prev >= 0;  // this segment starts at a cut
  int nfr_max = me.nfr;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) nfr_max = max(nfr_max, __shfl_xor_sync(0xffffffffu, nfr_max, o));
  if (nfr_max <= 0) return;

  // ---- interpolated coefficients c[k] = coefficients[k + 1] of this lane's segment, in registers
  // (mod.rs:116-125: cinc = (cc - coefficients) / fperiod; coefficients = mc2b of the frame before)
  const double* const brow = d.b + me.b_base + 1;
  auto frame_row = [&](int frame_rel) -> const double* {
    int f = me.warm + frame_rel;
    f = f < 0 ? 0 : (f > me.last_frame ? me.last_frame : f);
    return brow + (long long)f * kNM;
  };
#if JB_V3_CINC_SMEM
  double c[kTaps];
  double* const cincs = S.cinc + lane;
#define CINC(k) cincs[(k) * 32]
#else
  double c[kTaps], cinc[kTaps];
#define CINC(k) cinc[k]
#endif
  {
    // the frame before the first is frame 0 itself (mod.rs:79-109); with beta > 0 it is mc2b of the
    // un-filtered first frame
    const double* r0 = (me.warm == 0 && d.b_first != nullptr) ? d.b_first + (long long)info[g].utt * kNM + 1
                                                              : frame_row(-1);
    if (JB_V3_BINC) {
      const double* ri = d.binc + me.b_base + 1 + (long long)me.warm * kNM;
#pragma unroll
      for (int k = 0; k < kTaps; ++k) {
        c[k] = r0[k];
        CINC(k) = ri[k];
      }
    } else {
      const double* r1 = frame_row(0);
#pragma unroll
      for (int k = 0; k < kTaps; ++k) {
        c[k] = r0[k];
        CINC(k) = (r1[k] - c[k]) / fper;
      }
    }
  }

  // ---- filter state
  double t[kTaps];
#pragma unroll
  for (int i = 0; i < kTaps; ++i) t[i] = 0.0;
  double d11 = 0.0, gin = 0.0, uin = 0.0;

  // ---- input ring: chunks 0 and 1
  const int nsmp_max = nfr_max * fp;
  auto fill_chunk = [&](int chunk) {  // 6 slots x 16 samples = 96 values, 3 per lane
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int k = lane + 32 * r;
      const int gg = k >> 4, i = k & (kCH - 1);
      const SlotInfo& si = info[gg];
      const int n = chunk * kCH + i;
      double* dst = S.In + ((chunk & 1) * kCH + i) * 8 + gg;
      if (n < si.nfr * fp) cp_async8s(dst, d.wave + si.in_base + n);
      else *dst = 0.0;
    }
  };
#if JB_V3_ROWS_SMEM
  auto stage_row = [&](int frame_rel) {  // row `frame_rel` (relative to warm) of all six slots -> S.R[parity]
    double* const dst = S.R[frame_rel & 1];
#pragma unroll
    for (int r = 0; r < (kSlots * kTaps + 31) / 32; ++r) {
      const int idx = lane + 32 * r;
      if (idx < kSlots * kTaps) {
        const int gg = idx / kTaps, k = idx - kTaps * gg;
        const SlotInfo& si = info[gg];
        int f = si.warm + frame_rel;
        f = f < 0 ? 0 : (f > si.last_frame ? si.last_frame : f);
        cp_async8s(dst + k * 8 + gg, d.b + si.b_base + 1 + (long long)f * kNM + k);
      }
    }
  };
  stage_row(0);
  stage_row(1);
#endif
  fill_chunk(0);
  if (kCH < nsmp_max) fill_chunk(1);
  cp_async_wait_all_s();
  __syncwarp();

  auto dump_state = [&](int which) {  // which: 0 = top of the own range, 1 = end of the segment
    if (lane >= 30 || me.seg < 0 || sg.state == nullptr) return;
    double* o = sg.state + ((long long)me.seg * 2 + which) * kMlsaSegStateDoubles + s * 37;
#pragma unroll
    for (int i = 0; i < kTaps; ++i) o[i] = t[i];
    o[34] = d11;
    o[35] = gin;
    o[36] = uin;
  };

  int left = fp, frames_done = 0;
  int i0_start = 0;
  const double* inp = S.In + g;
#if JB_V3_ROWS_SMEM && JB_V3_F32_HEAD
  // ---- single-precision head of the warm-up. The output of the warm-up frames is discarded and what the
  // state keeps of an early frame decays by >= 0.2 decades per frame, so the first own0 - f64_tail warm-up
  // frames of a warp whose six segments all start at a cut run the same recurrence in FP32 (four times the
  // issue rate, half the chain latency); the rounding error it leaves (~1e-6 relative) fades below the cut
  // tolerance over the f64_tail double-precision frames that follow. Nothing is assumed here either: the cut
  // check compares the state at the top of the own range with the true one as before.
  int n32 = 0x7fffffff;
  if (lane < 30 && me.seg >= 0 && me.nfr > 0) n32 = max(0, me.own0 - f64_tail);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n32 = min(n32, __shfl_xor_sync(0xffffffffu, n32, o));
  if (f64_tail < 0 || n32 == 0x7fffffff) n32 = 0;
  if (n32 > 0) {
    float t32[kTaps], c32[kTaps], ci32[kTaps];
#pragma unroll
    for (int k = 0; k < kTaps; ++k) {
      t32[k] = 0.f;
      c32[k] = (float)c[k];
      ci32[k] = (float)CINC(k);
    }
    float d11s = 0.f, gins = 0.f, uins = 0.f;
    const float af = (float)a, maf = -af, iaaf = (float)iaa;
    const float q1 = (float)pp1, q2 = (float)pp2, q3 = (float)pp3, q4 = (float)pp4, q5 = (float)pp5;
    auto step32 = [&](float* __restrict__ Y, float* __restrict__ F, int jin) {
      float po = t32[0];
      float pn = __fmaf_rn(af, po, iaaf * uins);
      t32[0] = pn;
      float y0 = 0.f, y1 = 0.f;
#pragma unroll
      for (int i = 1; i < kTaps; i += 2) {
        float q = __fmaf_rn(af, t32[i], po);
        po = t32[i];
        pn = __fmaf_rn(maf, pn, q);
        t32[i] = pn;
        y0 = __fmaf_rn(c32[i], pn, y0);
        if (i + 1 < kTaps) {
          q = __fmaf_rn(af, t32[i + 1], po);
          po = t32[i + 1];
          pn = __fmaf_rn(maf, pn, q);
          t32[i + 1] = pn;
          y1 = __fmaf_rn(c32[i + 1], pn, y1);
        }
      }
      d11s = __fmaf_rn(af, d11s, iaaf * gins);
      Y[lane] = y0 + y1;
      F[lane] = d11s * c32[0];
      __syncwarp();
#pragma unroll
      for (int k = 0; k < kTaps; ++k) c32[k] += ci32[k];
      const float y1v = Y[g5], y2v = Y[g5 + 1], y3v = Y[g5 + 2], y4v = Y[g5 + 3], y5v = Y[g5 + 4];
      const float f1v = F[g5], f2v = F[g5 + 1], f3v = F[g5 + 2], f4v = F[g5 + 3], f5v = F[g5 + 4];
      const float yprev = Y[prev_lane], fprev = F[prev_lane];
      const float e = (float)inp[jin * 8];
      const float Af = __fmaf_rn(q5, f5v, __fmaf_rn(q3, f3v, q1 * f1v));
      const float Bf = __fmaf_rn(q4, f4v, q2 * f2v);
      const float Ay = __fmaf_rn(q5, y5v, __fmaf_rn(q3, y3v, q1 * y1v));
      const float By = __fmaf_rn(q4, y4v, q2 * y2v);
      const float d120 = (e + Af) - Bf;
      const float x1 = __fmaf_rn(2.f, Af, e);
      const float d220 = (x1 + Ay) - By;
      gins = s == 0 ? d120 : fprev;
      uins = s == 0 ? d220 : yprev;
    };
    float* const Y32 = reinterpret_cast<float*>(&S.Y[0][0]);  // [2][32] floats in the space of S.Y[0]
    float* const F32 = reinterpret_cast<float*>(&S.F[0][0]);
    const int n32_smp = n32 * fp;
    for (int i0 = 0; i0 < n32_smp; i0 += kCH) {
      const int ring = (i0 & kCH);
#pragma unroll 1
      for (int j = 0; j < kCH; j += 2) {
        step32(Y32, F32, ring + j);
        step32(Y32 + 32, F32 + 32, ring + j + 1);
      }
      left -= kCH;
      bool restage = false;
      if (left == 0) {  // frame end: rows from shared memory as in the double-precision part
        left = fp;
        if (frames_done + 1 < n32) {
          const double* r0 = S.R[frames_done & 1] + g;
          const double* r1 = S.R[(frames_done + 1) & 1] + g;
#pragma unroll
          for (int k = 0; k < kTaps; ++k) {
            const double cc = r0[k * 8];
            c32[k] = (float)cc;
            ci32[k] = (float)((r1[k * 8] - cc) / fper);
          }
          restage = true;
        }  // the last frame end hands over to double precision below, outside this loop
        ++frames_done;
      }
      cp_async_wait_all_s();
      __syncwarp();
      if ((i0 / kCH + 2) * kCH < nsmp_max) fill_chunk(i0 / kCH + 2);
      if (restage) stage_row(frames_done + 1);
      __syncwarp();
    }
    {  // hand-over: double-precision coefficients of the first FP64 frame (the frame end skipped above)
      const double* r0 = S.R[(n32 - 1) & 1] + g;
      const double* r1 = S.R[n32 & 1] + g;
#pragma unroll
      for (int k = 0; k < kTaps; ++k) {
        c[k] = r0[k * 8];
        CINC(k) = (r1[k * 8] - c[k]) / fper;
      }
      __syncwarp();
      stage_row(n32 + 1);  // waited for at the end of the next chunk
    }
#pragma unroll
    for (int k = 0; k < kTaps; ++k) t[k] = (double)t32[k];
    d11 = (double)d11s;
    gin = (double)gins;
    uin = (double)uins;
    i0_start = n32_smp;
  }
#endif

  bool staged = false;
  auto frame_end = [&]() {  // mod.rs:140: the coefficients restart from mc2b of the frame just filtered
    left = fp;
#if JB_V3_ROWS_SMEM
    const double* r0 = S.R[frames_done & 1] + g;
    const double* r1 = S.R[(frames_done + 1) & 1] + g;
    ++frames_done;
#pragma unroll
    for (int k = 0; k < kTaps; ++k) {
      c[k] = r0[k * 8];
      CINC(k) = (r1[k * 8] - c[k]) / fper;
    }
    staged = true;
#else
    const double* r0 = frame_row(frames_done);
    ++frames_done;
    if (JB_V3_BINC) {  // (cc - coefficients) / fperiod precomputed per frame (launch_row_increments)
      int f = me.warm + frames_done;
      const bool past = f > me.last_frame;  // beyond the utterance the coefficients stand still
      f = past ? me.last_frame : f;
      const double* ri = d.binc + me.b_base + 1 + (long long)f * kNM;
#pragma unroll
      for (int k = 0; k < kTaps; ++k) {
        c[k] = r0[k];
        CINC(k) = past ? 0.0 : ri[k];
      }
    } else {
      const double* r1 = frame_row(frames_done);
#pragma unroll
      for (int k = 0; k < kTaps; ++k) {
        c[k] = r0[k];
        CINC(k) = (r1[k] - c[k]) / fper;
      }
    }
#endif
  };

  auto step = [&](double* __restrict__ Y, double* __restrict__ F, int jin, int jout) {
    // ---------------- phase A: fir of this lane's stage (mlsa.rs:128-163) and its df1 section
    double po = t[0];
    double pn = __fma_rn(a, po, iaa * uin);
    t[0] = pn;
    double y0 = 0.0, y1 = 0.0;
#pragma unroll
    for (int i = 1; i < kTaps; i += 2) {  // taps d[i+1], d[i+2] with coefficients[i+1], [i+2]
      double q = __fma_rn(a, t[i], po);
      po = t[i];
      pn = __fma_rn(ma, pn, q);
      t[i] = pn;
      y0 = __fma_rn(c[i], pn, y0);
      if (i + 1 < kTaps) {
        q = __fma_rn(a, t[i + 1], po);
        po = t[i + 1];
        pn = __fma_rn(ma, pn, q);
        t[i + 1] = pn;
        y1 = __fma_rn(c[i + 1], pn, y1);
      }
    }
    d11 = __fma_rn(a, d11, iaa * gin);  // mlsa.rs:57-58
    Y[lane] = y0 + y1;
    F[lane] = d11 * c[0];
    __syncwarp();
    // coefficients of the next sample (mod.rs:133-135); independent of the exchange
#pragma unroll
    for (int k = 0; k < JB_V3_CADD_FRONT; ++k) c[k] += CINC(k);
#if JB_V3_BOUNDARY_IN_STEP
    if (--left == 0) frame_end();
#endif
    // ---------------- phase B: Pade combine of df1 then df2 (mlsa.rs:54-79), next stage inputs
    const double y1v = Y[g5], y2v = Y[g5 + 1], y3v = Y[g5 + 2], y4v = Y[g5 + 3], y5v = Y[g5 + 4];
    const double f1v = F[g5], f2v = F[g5 + 1], f3v = F[g5 + 2], f4v = F[g5 + 3], f5v = F[g5 + 4];
    const double yprev = Y[prev_lane], fprev = F[prev_lane];
    const double e = inp[jin * 8];
#pragma unroll
    for (int k = JB_V3_CADD_FRONT; k < kTaps; ++k) c[k] += CINC(k);  // these fill the latency of the loads
    const double Af = __fma_rn(pp5, f5v, __fma_rn(pp3, f3v, pp1 * f1v));
    const double Bf = __fma_rn(pp4, f4v, pp2 * f2v);
    const double Ay = __fma_rn(pp5, y5v, __fma_rn(pp3, y3v, pp1 * y1v));
    const double By = __fma_rn(pp4, y4v, pp2 * y2v);
    const double d120 = (e + Af) - Bf;        // d12[0]
    const double x1 = __fma_rn(2.0, Af, e);   // x after df1
    const double d220 = (x1 + Ay) - By;       // d22[0]
    const double outv = __fma_rn(2.0, Ay, x1);
    gin = s == 0 ? d120 : fprev;
    uin = s == 0 ? d220 : yprev;
    if (head) S.Out[jout * 8 + g] = outv;
  };

  for (int i0 = i0_start; i0 < nsmp_max; i0 += kCH) {
    if (left == fp) {  // top of a frame (fp % kCH == 0): state taps for the cut check
      if (frames_done == me.own0 && my_cut) dump_state(0);
    }
    const int ring = (i0 & kCH);  // which half of the input ring this chunk sits in
#pragma unroll 1
    for (int j = 0; j < kCH; j += 2) {
      step(S.Y[0], S.F[0], ring + j, j);
      step(S.Y[1], S.F[1], ring + j + 1, j + 1);
    }
#if !JB_V3_BOUNDARY_IN_STEP
    left -= kCH;  // fperiod % kCH == 0: a frame ends with a chunk
    if (left == 0) frame_end();
#endif
    if (left == fp && frames_done == me.nfr) dump_state(1);
    cp_async_wait_all_s();  // chunk i0/kCH + 1 has landed
    __syncwarp();
    {  // flush 6 x 16 outputs; refill the half of the ring that was just consumed
      const int chunk = i0 / kCH;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int k = lane + 32 * r;
        const int gg = k >> 4, i = k & (kCH - 1);
        const SlotInfo& si = info[gg];
        const int n = i0 + i;
        if (n >= si.own0 * fp && n < si.nfr * fp) d.out[si.in_base + n] = S.Out[i * 8 + gg] * p.volume;
      }
      if ((chunk + 2) * kCH < nsmp_max) fill_chunk(chunk + 2);
#if JB_V3_ROWS_SMEM
      if (staged) {
        staged = false;
        stage_row(frames_done + 1);
      }  // over the row just consumed (the __syncwarp above orders the reads)
#endif
    }
    __syncwarp();
  }
}

// Compares the two sides of every cut: state of segment g at the top of its own range (after its
// warm-up from zero) against the end state of the segment before it. max |difference| -> *worst.
// `relative` (one warp per block): the mismatch is divided by max(1, largest |state| of the true side).
// `prev_state`: the state array the `prev` ids index (nullptr: sg.state itself; a repair pass keeps its own
// states but is compared with the end states of the original pass). sg.mismatch[sid], when present,
// receives the mismatch of this segment's cut.
__global__ void k_mlsa_seg_check(MlsaSegments sg, const double* prev_state, double* worst, int stride, int count,
                                 int relative) {
  const int sid = blockIdx.x;
  const MlsaSegment s = sg.seg[sid];
  if (s.prev < 0) return;
  const double* A = sg.state + ((long long)sid * 2 + 0) * stride;
  const double* B = (prev_state ? prev_state : sg.state) + ((long long)s.prev * 2 + 1) * stride;
  double m = 0.0, scale = 1.0;
  for (int i = threadIdx.x; i < count; i += blockDim.x) {
    const double dd = fabs(A[i] - B[i]);
    m = (dd > m || dd != dd) ? (dd != dd ? __longlong_as_double(0x7ff0000000000000ll) : dd) : m;
    if (relative) scale = fmax(scale, fabs(B[i]));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (relative) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) scale = fmax(scale, __shfl_xor_sync(0xffffffffu, scale, o));
    m = m / scale;  // inf / inf = NaN is caught by the host's !(w <= tol)
  }
  if ((threadIdx.x & 31) == 0 && m > 0.0) {
    atomicMax(reinterpret_cast<unsigned long long*>(worst), (unsigned long long)__double_as_longlong(m));
    if (sg.mismatch)
      atomicMax(reinterpret_cast<unsigned long long*>(sg.mismatch + sid), (unsigned long long)__double_as_longlong(m));
  }
}

// The result goes to pinned host memory by a plain store from the device: a cudaMemcpy on this
// stream would queue in the copy engine behind the waveform copies of the copy stream and hold up
// everything launched after it.
__global__ void k_publish(const double* src, double* host_dst) {
  *host_dst = *src;
  __threadfence_system();
}

}  // namespace

int launch_mlsa_v3(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, const MlsaSegments& sg,
                   int seg_first, int seg_count, cudaStream_t stream, int f64_tail) {
  if (seg_count <= 0) return 0;
  if (p.nmcp != kNM || p.fperiod % kCH != 0 || d.out == nullptr || d.out == d.wave || seg_first < 0 ||
      (JB_V3_BINC && d.binc == nullptr) ||
      seg_first + seg_count > sg.n_seg)
    return (int)cudaErrorInvalidValue;
  const int warps = (seg_count + kSlots - 1) / kSlots;
  // 1 - alpha^2 as a parameter: derived in the kernel it was re-computed from the constant bank at the
  // top of every sample (254 registers leave no room to keep it): a load and two dependent FP64
  // operations on the serial path of the filter
  size_t smem = 0;
#if JB_V3_CINC_SMEM
  smem = sizeof(WarpShared) * kWPB;
  cudaError_t e = cudaFuncSetAttribute(k_mlsa_v3<kWPB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
#endif
  k_mlsa_v3<kWPB><<<(warps + kWPB - 1) / kWPB, 32 * kWPB, smem, stream>>>(b, p, d, sg, seg_first,
                                                                        seg_first + seg_count, 1.0 - p.alpha * p.alpha, f64_tail);
  return (int)cudaGetLastError();
}

int launch_mlsa_seg_check(const MlsaSegments& sg, double* worst, double* host_mirror, cudaStream_t stream,
                          int stride, int count, bool relative, const double* prev_state) {
  if (sg.n_seg == 0 || worst == nullptr || sg.state == nullptr) return 0;
  cudaError_t e = cudaMemsetAsync(worst, 0, sizeof(double), stream);
  if (e == cudaSuccess && sg.mismatch) e = cudaMemsetAsync(sg.mismatch, 0, (size_t)sg.n_seg * sizeof(double), stream);
  if (e != cudaSuccess) return (int)e;
  k_mlsa_seg_check<<<sg.n_seg, relative ? 32 : 64, 0, stream>>>(sg, prev_state, worst, stride, count, relative ? 1 : 0);
  if (host_mirror) k_publish<<<1, 1, 0, stream>>>(worst, host_mirror);
  return (int)cudaGetLastError();
}

}  // namespace jbk
