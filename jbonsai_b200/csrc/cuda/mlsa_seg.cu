// MLSA filter, variant 3 (default): time-segmented, six segments per warp, one lane per Pade stage.
//
// Reference: MelLogSpectrumApproximation::<6>::{df, df1, df2, fir} (src/vocoder/mlsa.rs:48-163)
// driven per sample by Vocoder::synthesize (src/vocoder/mod.rs:112-141).
//
// Why segments. The filter is a feedback loop along time, so one utterance is one serial chain
// and a batch of 1024 utterances can keep at most 1024 chains busy on a chip with 592 warp
// schedulers: variants 0-2 are latency-bound. The filter's memory fades, though: started from a zero
// state W frames early, its state is indistinguishable from the true one (to the rounding-noise
// floor) after a few tens of frames. The per-sample coefficients need no history at all (they restart
// from mc2b of the previous frame at every frame boundary, mod.rs:140). So an utterance is cut into
// segments; every segment but the first runs `warm` extra frames in front of its own range with its
// output discarded. Nothing is assumed: every segment stores its state at the start of its own range
// and at its end, k_mlsa_seg_check compares the two sides of every cut and lists the cuts whose
// mismatch could move the waveform by more than the bar; the host re-filters the head of those
// segments from the exact state (launch_mlsa_v3 with `init`), see Engine::mlsa_repair_segments.
//
// Layout. lane = 5 g + s: segment slot g = 0..5 of this warp, Pade stage s + 1 (s = 0..4); lanes
// 30/31 shadow slot 5 and write to dump slots. A lane owns the whole warped delay line of its stage
// in registers and runs the tap recurrence serially in the "q form"
//   t'_1 = (1-a^2) u + a t_1 ;  t'_j = (t_{j-1} + a t_j) - a t'_{j-1}   (== iaa*rem + a*d of mlsa.rs:141-153)
// with two dot-product accumulators like the reference (y0/y1): three FMAs per tap, the arithmetic
// floor of the recurrence. Every lane also keeps the interpolated coefficients of its segment and
// their per-sample increments in registers and advances them itself (mod.rs:133-135): sharing them
// between the five stage lanes through shared memory was measured slower (tools/ubench4.cu: broadcast
// LDS.128 of the coefficients costs more than the 27 additions it saves). One shared-memory exchange
// per sample carries the five stage outputs and the five df1 outputs to the lanes of the segment; the
// Pade combine (d[0] = (x + A) - B, x' = x + 2A with A/B the odd/even weighted sums; mlsa.rs:54-79)
// is done redundantly by each lane.
//
// Loop structure (round 2). frames { prologue: coefficients and increments of the frame from shared
// memory ; chunks of 16 samples { sample steps ; flush 16 outputs, refill the input ring } }. The
// frame boundary used to sit inside the sample step (a divergent block with 34 inlined divisions, once
// per 240 samples): the step is now straight-line code, the increments (cc - c) / fperiod come
// precomputed per frame (launch_row_increments) and are staged with the rows by cp.async one frame
// ahead.
//
// Cost model (tools/ubench3.cu, ubench4.cu on B200): an FP64 instruction occupies a scheduler's FP64
// pipe for max(2, number of distinct 64-bit register sources) cycles (uniform-register and constant
// operands are free), a dependent DFMA issues 8 cycles after its producer. Per tap: q (2) + chain (2) +
// dot product (3) + coefficient add (2) = 9 cycles against a chain latency of 8, so two warps per
// scheduler keep the pipe busy and a third cannot add anything.
#include <type_traits>

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

constexpr int kSlots = 6;     // segments per warp
constexpr int kCHMax = 80;    // samples per staged chunk: the largest even divisor of fperiod up to this (240 -> 80)
constexpr int kWPB = 4;       // warps per block (independent of each other)
constexpr int kInStride = 2 * kCHMax + 1;  // odd strides: the six slots of a sample index sit in six different banks
constexpr int kOutStride = kCHMax + 1;

struct SlotInfo {
  long long in_base;   // index of the first sample this segment filters (frame `warm`)
  long long b_base;    // frame_off[utt] * NM
  int warm;            // first filtered frame
  int last_frame;      // utterance length - 1
  int nfr;             // frames filtered = end - warm (0: empty slot)
  int own0;            // begin - warm: frames whose output is discarded
  int seg;             // segment id (state slot), -1 for an empty slot
  int prev;            // id of the segment before this one in the utterance, -1 if none
  int utt;             // utterance (row of VocoderDev::b_first)
  int init;            // 1: start from the end state of `prev` in MlsaSegments::init_state (no warm-up)
};

// Shared memory of one warp. Staging arrays are [slot][index] with odd row strides: the copies in and out
// walk along a row (consecutive banks), the filter lanes read one index of all six rows (six different banks).
template <int NM>
struct WarpShared {
  double X[2][128];                 // per-sample exchange, double-buffered by sample parity: stage outputs y and df1
                                    // section outputs f of every slot by lane, raw and weighted (see step)
  double In[kSlots * kInStride];    // excitation ring: [slot][sample mod 2 CH]
  double Out[kSlots * kOutStride];  // output tile: [slot][sample mod CH]
  double Rc[2][kSlots * NM];        // coefficients a frame starts from (mc2b row of the frame before), [frame parity][slot][tap]
  double Ri[2][kSlots * NM];        // their per-sample increments over that frame
  SlotInfo info[kSlots];
};

template <int NM, int WPB>
__global__ void __launch_bounds__(32 * WPB, 2)
k_mlsa_v3(BatchLayout b, VocoderParams p, VocoderDev d, MlsaSegments sg, int seg_first, int seg_end, const double iaa,
          const int CH) {
  constexpr int kTaps = NM - 1;  // delay-line taps d[1..NM-1] (d[0] is write-only in the reference)
  extern __shared__ __align__(16) unsigned char v3_smem[];  // WarpShared<NM>[WPB]: above the 48 KB static limit
  WarpShared<NM>* const sh = reinterpret_cast<WarpShared<NM>*>(v3_smem);

  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int warp = blockIdx.x * WPB + wib;
  if (seg_first + warp * kSlots >= seg_end) return;  // whole warp leaves; warps of a block never sync with each other
  WarpShared<NM>& S = sh[wib];
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
      si.b_base = foff * NM;
      si.warm = s.warm;
      si.last_frame = (int)(b.frame_off[s.utt + 1] - foff) - 1;
      si.nfr = s.end - s.warm;
      si.own0 = s.begin - s.warm;
      si.seg = sid;
      si.prev = s.prev;
      si.utt = s.utt;
      si.init = (sg.init_state != nullptr && s.prev >= 0 && s.warm == s.begin) ? 1 : 0;
    } else {
      si.in_base = 0; si.b_base = 0; si.warm = 0; si.last_frame = 0; si.nfr = 0; si.own0 = 0; si.seg = -1; si.prev = -1;
      si.utt = 0; si.init = 0;
    }
    info[lane] = si;
  }
  __syncwarp();

  const int g = lane < 30 ? lane / 5 : 5, s = lane < 30 ? lane - 5 * g : lane - 30;
  const int g5 = 5 * g;
  const int prev_lane = s > 0 ? g5 + s - 1 : g5;
  const bool head = lane < 30 && s == 0;
  const double a = p.alpha, ma = -a;  // iaa = 1 - a^2 is a kernel parameter (see the launcher)
  // Per-lane scalars that are only needed at frame and chunk boundaries stay in shared memory (info[g]) and
  // the loop counters are warp-uniform (REDUX -> uniform registers): the sample step itself needs every one of
  // the 255 vector registers (three arrays of NM - 1 doubles are 204 of them).
  const int nfr_max = __reduce_max_sync(0xffffffffu, info[g].nfr);
  if (nfr_max <= 0) return;
  // Pade weight of this lane's stage: outputs are exchanged raw (next stage's input) and weighted (the combine)
  const double ppw = c_pp3[s + 1];

  // ---- staging: frame `frame_rel` (relative to warm) of all six slots -> S.Rc / S.Ri[parity], a row per pass.
  // Coefficients at the top of frame F = mc2b row of frame F - 1 (mod.rs:140); the frame before the first
  // is frame 0 itself (mod.rs:79-109) or, with beta > 0, mc2b of the un-filtered first frame (b_first).
  // Increments (cc - coefficients) / fperiod: VocoderDev::binc. Past the end of the utterance (a shorter
  // segment idling beside a longer one) the coefficients stand still.
  auto stage_row = [&](int frame_rel) {
    double* const dc = S.Rc[frame_rel & 1];
    double* const di = S.Ri[frame_rel & 1];
#pragma unroll 1
    for (int gg = 0; gg < kSlots; ++gg) {
      const SlotInfo& si = info[gg];
      const int f = si.warm + frame_rel;  // warp-uniform
      const int fc = f - 1 > si.last_frame ? si.last_frame : f - 1;
      const double* src = (fc < 0) ? (d.b_first != nullptr ? d.b_first + (long long)si.utt * NM + 1 : d.b + si.b_base + 1)
                                   : d.b + si.b_base + 1 + (long long)fc * NM;
      const double* inc = d.binc + si.b_base + 1 + (long long)f * NM;
      const bool past = f > si.last_frame;
      for (int k = lane; k < kTaps; k += 32) {
        cp_async8s(dc + gg * NM + k, src + k);
        if (past) di[gg * NM + k] = 0.0;
        else cp_async8s(di + gg * NM + k, inc + k);
      }
    }
  };

  // ---- input ring: chunk `chunk` of all six slots -> its half of S.In
  const int nsmp_max = nfr_max * fp;
  auto fill_chunk = [&](int chunk) {
    const int half = (chunk & 1) * CH;
    const int n0 = chunk * CH;
#pragma unroll 1
    for (int gg = 0; gg < kSlots; ++gg) {
      const SlotInfo& si = info[gg];
      const int nend = si.nfr * fp;
      const double* src = d.wave + si.in_base + n0;
      double* dst = S.In + gg * kInStride + half;
      for (int i = lane; i < CH; i += 32) {
        if (n0 + i < nend) cp_async8s(dst + i, src + i);
        else dst[i] = 0.0;
      }
    }
  };
  stage_row(0);
  fill_chunk(0);
  if (CH < nsmp_max) fill_chunk(1);

  // ---- filter state: zero, or the exact end state of the segment before (repair pass)
  double t[kTaps];
  double d11 = 0.0, gin = 0.0, uin = 0.0;
  if (info[g].init && lane < 30) {
    const double* o = sg.init_state + ((long long)info[g].prev * 2 + 1) * kMlsaSegStateDoubles + s * 37;
#pragma unroll
    for (int i = 0; i < kTaps; ++i) t[i] = o[i];
    d11 = o[34];
    gin = o[35];
    uin = o[36];
  } else {
#pragma unroll
    for (int i = 0; i < kTaps; ++i) t[i] = 0.0;
  }
  cp_async_wait_all_s();
  __syncwarp();

  auto dump_state = [&](int which) {  // which: 0 = top of the own range, 1 = end of the segment
    const int seg = info[g].seg;
    if (lane >= 30 || seg < 0 || sg.state == nullptr) return;
    double* o = sg.state + ((long long)seg * 2 + which) * kMlsaSegStateDoubles + s * 37;
#pragma unroll
    for (int i = 0; i < kTaps; ++i) o[i] = t[i];
#pragma unroll
    for (int i = kTaps; i < 34; ++i) o[i] = 0.0;
    o[34] = d11;
    o[35] = gin;
    o[36] = uin;
  };

  double c[kTaps], cinc[kTaps];
  const double* const in_row = S.In + g * kInStride;
  double* const out_row = S.Out + g * kOutStride;

  // Exchange buffers of one sample parity: X[0..31] raw stage outputs y, X[32..63] raw df1 outputs f,
  // X[64..95] / X[96..127] the same weighted by the Pade coefficient of the stage.
  auto step = [&](double* __restrict__ X, int jin, int jout) {
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
    {
      const double y = y0 + y1, fo = d11 * c[0];
      X[lane] = y;
      X[32 + lane] = fo;
      X[64 + lane] = ppw * y;
      X[96 + lane] = ppw * fo;
    }
    __syncwarp();
    // ---------------- phase B: Pade combine of df1 then df2 (mlsa.rs:54-79), next stage inputs.
    // A = weighted sum over the odd stages, B over the even ones: d[0] = (x + A) - B, x' = x + 2 A.
    const double* W = X + 64 + g5;
    const double wy1 = W[0], wy2 = W[1], wy3 = W[2], wy4 = W[3], wy5 = W[4];
    const double wf1 = W[32], wf2 = W[33], wf3 = W[34], wf4 = W[35], wf5 = W[36];
    const double yprev = X[prev_lane], fprev = X[32 + prev_lane];
    const double e = in_row[jin];
    // coefficients of the next sample (mod.rs:133-135); independent of the exchange: they fill the latency of the loads
#pragma unroll
    for (int k = 0; k < kTaps; ++k) c[k] += cinc[k];
    const double Af = wf5 + (wf3 + wf1);
    const double Bf = wf4 + wf2;
    const double Ay = wy5 + (wy3 + wy1);
    const double By = wy4 + wy2;
    const double d120 = (e + Af) - Bf;        // d12[0]
    const double x1 = __fma_rn(2.0, Af, e);   // x after df1
    const double d220 = (x1 + Ay) - By;       // d22[0]
    const double outv = __fma_rn(2.0, Ay, x1);
    gin = s == 0 ? d120 : fprev;
    uin = s == 0 ? d220 : yprev;
    if (head) out_row[jout] = outv;
  };

  const int chunks_per_frame = fp / CH;  // fperiod % CH == 0, CH even (launcher)
  int chunk = 0;
  for (int f = 0; f < nfr_max; ++f) {
    {  // frame prologue (mod.rs:116-125,140)
      const double* rc = S.Rc[f & 1] + g * NM;
      const double* ri = S.Ri[f & 1] + g * NM;
#pragma unroll
      for (int k = 0; k < kTaps; ++k) {
        c[k] = rc[k];
        cinc[k] = ri[k];
      }
    }
    if (f + 1 < nfr_max) stage_row(f + 1);  // the other parity: last read in the prologue of frame f - 1
    if (f == info[g].own0 && info[g].prev >= 0 && !info[g].init) dump_state(0);  // top of the own range behind a cut
    for (int cf = 0; cf < chunks_per_frame; ++cf, ++chunk) {
      const int ring = (chunk & 1) * CH;  // which half of the input ring this chunk sits in
#pragma unroll 1
      for (int j = 0; j < CH; j += 2) {
        step(S.X[0], ring + j, j);
        step(S.X[1], ring + j + 1, j + 1);
      }
      cp_async_wait_all_s();  // chunk + 1 (and the next frame's rows) have landed
      __syncwarp();
      {  // flush 6 x CH outputs; refill the half of the ring that was just consumed
        const int n0 = chunk * CH;
#pragma unroll 1
        for (int gg = 0; gg < kSlots; ++gg) {
          const SlotInfo& si = info[gg];
          const int nlo = si.own0 * fp - n0, nhi = si.nfr * fp - n0;  // own samples of this chunk: [nlo, nhi)
          double* dst = d.out + si.in_base + n0;
          const double* src = S.Out + gg * kOutStride;
          for (int i = lane; i < CH; i += 32)
            if (i >= nlo && i < nhi) dst[i] = src[i] * p.volume;
        }
        if ((chunk + 2) * CH < nsmp_max) fill_chunk(chunk + 2);
      }
      __syncwarp();
    }
    if (f + 1 == info[g].nfr) dump_state(1);
  }
}

// Compares the two sides of every cut: state of segment g at the top of its own range (after its
// warm-up from zero) against the end state of the segment before it. max |difference| -> *worst.
// `relative` (one warp per block): the mismatch is divided by max(1, largest |state value| of the true side).
// `prev_state`: the state array the `prev` ids index (nullptr: sg.state itself; a repair pass keeps its own
// states but is compared with the end states of the original pass). sg.mismatch[sid], when present,
// receives the mismatch of this segment's cut.
// `gain` (MLSA, stage 0): what a state error can do to the waveform. A tap error e moves a stage output by at
// most e * sum_k |c_k| and the waveform by about as much (the Pade weights of the odd stages sum to ~0.51 and
// enter twice), times the volume: a cut fails when mismatch * max(1, sum_k |c_k| of the frame at the cut) *
// volume > tol, i.e. `tol` is a bound on the waveform error (default 5e-7, half the 1e-6 bar). Failing
// segments are appended to sg.fail_list (count in sg.fail_count).
__global__ void k_mlsa_seg_check(MlsaSegments sg, const double* prev_state, double* worst, int stride, int count,
                                 int relative, BatchLayout b, const double* rows, int nm, double volume, double tol) {
  const int sid = blockIdx.x;
  const MlsaSegment s = sg.seg[sid];
  if (s.prev < 0) return;
  const double* A = sg.state + ((long long)sid * 2 + 0) * stride;
  const double* B = (prev_state ? prev_state : sg.state) + ((long long)s.prev * 2 + 1) * stride;
  double m = 0.0, scale = 1.0, gain = 0.0;
  for (int i = threadIdx.x; i < count; i += blockDim.x) {
    const double dd = fabs(A[i] - B[i]);
    m = (dd > m || dd != dd) ? (dd != dd ? __longlong_as_double(0x7ff0000000000000ll) : dd) : m;
    if (relative) scale = fmax(scale, fabs(B[i]));
  }
  if (rows != nullptr) {  // sum |c_k| of the rows on both sides of the cut (the coefficients move between them)
    const long long base = b.frame_off[s.utt] * nm;
    const int f0 = s.begin > 0 ? s.begin - 1 : 0;
    for (int i = threadIdx.x; i < 2 * nm; i += blockDim.x) {
      const int k = i % nm;
      const int f = i < nm ? f0 : s.begin;
      if (k >= 1) gain += fabs(rows[base + (long long)f * nm + k]);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    gain += __shfl_xor_sync(0xffffffffu, gain, o);
  }
  if (relative) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) scale = fmax(scale, __shfl_xor_sync(0xffffffffu, scale, o));
    m = m / scale;  // inf / inf = NaN is caught by the host's !(w <= tol)
  }
  __shared__ double sm[2], sgn[2];
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { sm[w] = m; sgn[w] = gain; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < (int)(blockDim.x >> 5); ++i) { m = fmax(m, sm[i]); gain += sgn[i]; }
    if (rows != nullptr) m = m * fmax(1.0, 0.5 * gain) * fabs(volume);  // 0.5: the two rows were summed
    if (m > 0.0) {
      atomicMax(reinterpret_cast<unsigned long long*>(worst), (unsigned long long)__double_as_longlong(m));
      if (sg.mismatch) sg.mismatch[sid] = m;
    }
    if (!(m <= tol) && sg.fail_count != nullptr) {
      const int slot = atomicAdd(sg.fail_count, 1);
      if (slot < sg.fail_cap) sg.fail_list[slot] = sid;
    }
  }
}

// The result goes to pinned host memory by a plain store from the device: a cudaMemcpy on this
// stream would queue in the copy engine behind the waveform copies of the copy stream and hold up
// everything launched after it.
__global__ void k_publish(const double* src, double* host_dst, const int* fail_count, int* host_fail) {
  *host_dst = *src;
  if (fail_count != nullptr && host_fail != nullptr) *host_fail = *fail_count;
  __threadfence_system();
}

// Repair pass epilogue: segment i of `fix` re-filtered frames [begin, end) of its utterance from the exact
// state into `scratch` (same indexing as `out`). When the last frame of the new samples agrees with the
// original pass (the two trajectories have merged: the rest of the original segment stands) the new samples
// replace the old ones; otherwise ok[i] = 0 and nothing is written.
__global__ void k_mlsa_repair_merge(BatchLayout b, const MlsaSegment* fix, int n_fix, const double* scratch, double* out,
                                    int fperiod, double tol, int* ok) {
  const int i = blockIdx.x;
  if (i >= n_fix) return;
  const MlsaSegment s = fix[i];
  const long long base = b.frame_off[s.utt] * (long long)fperiod;
  const long long n0 = base + (long long)s.begin * fperiod, n1 = base + (long long)s.end * fperiod;
  const long long utt_end = b.frame_off[s.utt + 1] * (long long)fperiod;
  __shared__ int bad;
  if (threadIdx.x == 0) bad = 0;
  __syncthreads();
  if (n1 < utt_end) {  // a repair range that reaches the end of the utterance has nothing to merge with
    for (long long n = n1 - fperiod + threadIdx.x; n < n1; n += blockDim.x)
      if (!(fabs(scratch[n] - out[n]) <= tol)) bad = 1;
  }
  __syncthreads();
  if (bad) {
    if (threadIdx.x == 0) ok[i] = 0;
    return;
  }
  for (long long n = n0 + threadIdx.x; n < n1; n += blockDim.x) out[n] = scratch[n];
  if (threadIdx.x == 0) ok[i] = 1;
}

}  // namespace

bool mlsa_v3_supports(int nmcp) { return nmcp == 35 || nmcp == 25; }

// Samples per staged chunk: the largest even divisor of fperiod that fits the staging buffers (0: none, odd fperiod).
int mlsa_v3_chunk(int fperiod) {
  for (int ch = kCHMax; ch >= 2; ch -= 2)
    if (fperiod % ch == 0) return ch;
  return 0;
}

template <int NM>
static int launch_v3_t(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, const MlsaSegments& sg,
                       int seg_first, int seg_count, cudaStream_t stream) {
  const int warps = (seg_count + kSlots - 1) / kSlots;
  // 1 - alpha^2 as a parameter: derived in the kernel it was re-computed from the constant bank at the
  // top of every sample: a load and two dependent FP64 operations on the serial path of the filter
  const size_t smem = sizeof(WarpShared<NM>) * kWPB;
  {  // per device (context), not per process: set it at every launch, it costs microseconds
    const cudaError_t e = cudaFuncSetAttribute(k_mlsa_v3<NM, kWPB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  k_mlsa_v3<NM, kWPB><<<(warps + kWPB - 1) / kWPB, 32 * kWPB, smem, stream>>>(
      b, p, d, sg, seg_first, seg_first + seg_count, 1.0 - p.alpha * p.alpha, mlsa_v3_chunk(p.fperiod));
  return (int)cudaGetLastError();
}

int launch_mlsa_v3(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, const MlsaSegments& sg,
                   int seg_first, int seg_count, cudaStream_t stream) {
  if (seg_count <= 0) return 0;
  if (!mlsa_v3_supports(p.nmcp) || mlsa_v3_chunk(p.fperiod) == 0 || d.out == nullptr || d.out == d.wave || seg_first < 0 ||
      d.binc == nullptr || seg_first + seg_count > sg.n_seg)
    return (int)cudaErrorInvalidValue;
  if (p.nmcp == 25) return launch_v3_t<25>(b, p, d, sg, seg_first, seg_count, stream);
  return launch_v3_t<35>(b, p, d, sg, seg_first, seg_count, stream);
}

int launch_mlsa_seg_check(const MlsaSegments& sg, double* worst, double* host_mirror, cudaStream_t stream,
                          int stride, int count, bool relative, const double* prev_state, const SegCheckGain* gain) {
  if (sg.n_seg == 0 || worst == nullptr || sg.state == nullptr) return 0;
  cudaError_t e = cudaMemsetAsync(worst, 0, sizeof(double), stream);
  if (e == cudaSuccess && sg.mismatch) e = cudaMemsetAsync(sg.mismatch, 0, (size_t)sg.n_seg * sizeof(double), stream);
  if (e == cudaSuccess && sg.fail_count) e = cudaMemsetAsync(sg.fail_count, 0, sizeof(int), stream);
  if (e != cudaSuccess) return (int)e;
  BatchLayout bl{};
  if (gain) bl = *gain->layout;
  k_mlsa_seg_check<<<sg.n_seg, relative ? 32 : 64, 0, stream>>>(
      sg, prev_state, worst, stride, count, relative ? 1 : 0, bl, gain ? gain->rows : nullptr, gain ? gain->nm : 0,
      gain ? gain->volume : 1.0, gain ? gain->tol : 0.0);
  if (host_mirror) k_publish<<<1, 1, 0, stream>>>(worst, host_mirror, sg.fail_count, sg.host_fail_count);
  return (int)cudaGetLastError();
}

int launch_mlsa_repair_merge(const BatchLayout& b, const MlsaSegment* fix, int n_fix, const double* scratch, double* out,
                             int fperiod, double tol, int* ok, cudaStream_t stream) {
  if (n_fix <= 0) return 0;
  k_mlsa_repair_merge<<<n_fix, 256, 0, stream>>>(b, fix, n_fix, scratch, out, fperiod, tol, ok);
  return (int)cudaGetLastError();
}

}  // namespace jbk
