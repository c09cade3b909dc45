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
// and the host re-runs the sequential kernel (variant 2) when a cut disagrees by more than 1e-8.
//
// Layout. lane = 5 g + s: segment slot g = 0..5 of this warp, Pade stage s + 1 (s = 0..4); lanes
// 30/31 shadow slot 5 and write to dump slots. A lane owns the whole 34-tap warped delay line of
// its stage in registers and runs the tap recurrence serially in the "q form"
//   t'_1 = (1-a^2) u + a t_1 ;  t'_j = (t_{j-1} + a t_j) - a t'_{j-1}   (== iaa*rem + a*d of mlsa.rs:141-153)
// with two dot-product accumulators like the reference (y0/y1). No scan, no carry exchange: three
// FMAs per tap, which is the arithmetic floor of the recurrence. The 35 interpolated coefficients
// of a segment live in shared memory (double-buffered by sample parity) and are advanced by the
// whole warp, 7 values per lane, so that the c += cinc additions (mod.rs:133-135) are not repeated
// by each of the five stage lanes. One shared-memory exchange per sample carries the five stage
// outputs and the five df1 outputs to the lanes of the segment; the Pade combine
// (d[0] = (x + A) - B, x' = x + 2A with A/B the odd/even weighted sums; mlsa.rs:54-79) is done
// redundantly by each lane.
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

constexpr int kNM = 35;       // mel-cepstral order + 1
constexpr int kTaps = 34;     // delay-line taps d[1..34] (d[0] is write-only in the reference)
constexpr int kSlots = 6;     // segments per warp
constexpr int kCH = 16;       // samples per staged chunk (input ring of 2 chunks, output tile of 1)
constexpr int kCRow = 38;     // doubles per coefficient row: 16-byte reads of 6 rows hit 6 bank groups
constexpr int kCBuf = kSlots * kCRow + 32;  // + dump slots for idle updater entries
constexpr int kXRow = 16;     // exchange row: y_1..y_5 at 0..4, f_1..f_5 at 8..12
constexpr int kXBuf = (kSlots + 1) * kXRow;  // + one dump row (lanes 30/31)
constexpr int kUpd = 7;       // coefficient entries advanced per lane: 6 * 34 = 204 <= 7 * 32
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
};

template <int WPB>
__global__ void __launch_bounds__(32 * WPB)
k_mlsa_v3(BatchLayout b, VocoderParams p, VocoderDev d, MlsaSegments sg) {
  __shared__ __align__(16) double sC[WPB][2][kCBuf];
  __shared__ __align__(16) double sX[WPB][2][kXBuf];
  __shared__ __align__(16) double sIn[WPB][kSlots * 2 * kCH];
  __shared__ __align__(16) double sOut[WPB][kSlots * kCH];
  __shared__ SlotInfo sInfo[WPB][kSlots];

  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int warp = blockIdx.x * WPB + wib;
  if (warp * kSlots >= sg.n_seg) return;  // whole warp leaves; warps of a block never sync with each other
  double* const C0 = sC[wib][0];
  double* const C1 = sC[wib][1];
  double* const X0 = sX[wib][0];
  double* const X1 = sX[wib][1];
  double* const In = sIn[wib];
  double* const Out = sOut[wib];
  SlotInfo* const info = sInfo[wib];

  const int fp = p.fperiod;
  // ---- slot table
  if (lane < kSlots) {
    SlotInfo si;
    const int sid = warp * kSlots + lane;
    if (sid < sg.n_seg) {
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
    } else {
      si.in_base = 0; si.b_base = 0; si.warm = 0; si.last_frame = 0; si.nfr = 0; si.own0 = 0; si.seg = -1; si.prev = -1;
    }
    info[lane] = si;
  }
  for (int i = lane; i < kCBuf; i += 32) { C0[i] = 0.0; C1[i] = 0.0; }
  for (int i = lane; i < kXBuf; i += 32) { X0[i] = 0.0; X1[i] = 0.0; }
  __syncwarp();

  const int g = lane < 30 ? lane / 5 : 5, s = lane < 30 ? lane - 5 * g : lane - 30;
  const int gx = lane < 30 ? g : kSlots;           // exchange row this lane writes (dump row for 30/31)
  const int sm1 = s > 0 ? s - 1 : 0;
  const bool head = lane < 30 && s == 0;
  const double a = p.alpha, ma = -a, iaa = 1.0 - a * a;
  const double pp1 = c_pp3[1], pp2 = c_pp3[2], pp3 = c_pp3[3], pp4 = c_pp3[4], pp5 = c_pp3[5];
  const double fper = (double)fp;
  const int my_nfr = info[g].nfr, my_own0 = info[g].own0, my_seg = info[g].seg;
  const bool my_cut = info[g].prev >= 0;  // this segment starts at a cut
  int nfr_max = my_nfr;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) nfr_max = max(nfr_max, __shfl_xor_sync(0xffffffffu, nfr_max, o));
  if (nfr_max <= 0) return;

  // ---- coefficient entries advanced by this lane: entry k = lane + 32 m -> (slot k / 34, c[1 + k % 34])
  int uaddr[kUpd];          // offset inside a coefficient buffer
  double creg[kUpd], ctgt[kUpd], cinc[kUpd];
  auto load_b = [&](int m, int frame_rel) -> double {
    const int k = lane + 32 * m;
    if (k >= kSlots * kTaps) return 0.0;
    const SlotInfo& si = info[k / kTaps];
    if (si.nfr <= 0) return 0.0;
    int f = si.warm + frame_rel;
    f = f < 0 ? 0 : (f > si.last_frame ? si.last_frame : f);
    return d.b[si.b_base + (long long)f * kNM + 1 + k % kTaps];
  };
#pragma unroll
  for (int m = 0; m < kUpd; ++m) {
    const int k = lane + 32 * m;
    uaddr[m] = k < kSlots * kTaps ? (k / kTaps) * kCRow + 1 + k % kTaps : kSlots * kCRow + lane;
    creg[m] = load_b(m, -1);   // coefficients at the top of frame `warm`: mc2b of the frame before
    ctgt[m] = load_b(m, 0);    // (of frame 0 itself for the first frame, mod.rs:79-109)
    cinc[m] = (ctgt[m] - creg[m]) / fper;
    C0[uaddr[m]] = creg[m];
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
      double* dst = In + gg * 2 * kCH + (chunk & 1) * kCH + i;
      if (n < si.nfr * fp) cp_async8s(dst, d.wave + si.in_base + n);
      else *dst = 0.0;
    }
  };
  fill_chunk(0);
  if (kCH < nsmp_max) fill_chunk(1);
  cp_async_wait_all_s();
  __syncwarp();

  auto dump_state = [&](int which) {  // which: 0 = top of the own range, 1 = end of the segment
    if (lane >= 30 || my_seg < 0 || sg.state == nullptr) return;
    double* o = sg.state + ((long long)my_seg * 2 + which) * kMlsaSegStateDoubles + s * 37;
#pragma unroll
    for (int i = 0; i < kTaps; ++i) o[i] = t[i];
    o[34] = d11;
    o[35] = gin;
    o[36] = uin;
  };

  const double* const Cc0 = C0 + g * kCRow;
  const double* const Cc1 = C1 + g * kCRow;
  const int wy = gx * kXRow + s, wf = gx * kXRow + 8 + s;
  const int rx = g * kXRow;
  int left = fp, frames_done = 0, inIdx = 0;

  auto step = [&](const double* __restrict__ Cc, double* __restrict__ Cn, double* __restrict__ X, int jout) {
    // ---------------- phase A: fir of this lane's stage (mlsa.rs:128-163) and its df1 section
    const double c1 = Cc[1];
    double po = t[0];
    double pn = __fma_rn(a, po, iaa * uin);
    t[0] = pn;
    double y0 = 0.0, y1 = 0.0;
#pragma unroll
    for (int i = 1; i < kTaps; i += 2) {  // taps j = i + 1 and i + 2 with coefficients c[j], c[j+1]
      const double2 cj = *reinterpret_cast<const double2*>(Cc + i + 1);
      double q = __fma_rn(a, t[i], po);
      po = t[i];
      pn = __fma_rn(ma, pn, q);
      t[i] = pn;
      y0 = __fma_rn(cj.x, pn, y0);
      if (i + 1 < kTaps) {
        q = __fma_rn(a, t[i + 1], po);
        po = t[i + 1];
        pn = __fma_rn(ma, pn, q);
        t[i + 1] = pn;
        y1 = __fma_rn(cj.y, pn, y1);
      }
    }
    d11 = __fma_rn(a, d11, iaa * gin);  // mlsa.rs:57-58
    X[wy] = y0 + y1;
    X[wf] = d11 * c1;
    // coefficients of the next sample (mod.rs:133-135, :140)
#pragma unroll
    for (int m = 0; m < kUpd; ++m) creg[m] += cinc[m];
    if (--left == 0) {
      left = fp;
      ++frames_done;
#pragma unroll
      for (int m = 0; m < kUpd; ++m) {
        creg[m] = ctgt[m];
        ctgt[m] = load_b(m, frames_done);
        cinc[m] = (ctgt[m] - creg[m]) / fper;
      }
    }
#pragma unroll
    for (int m = 0; m < kUpd; ++m) Cn[uaddr[m]] = creg[m];
    __syncwarp();
    // ---------------- phase B: Pade combine of df1 then df2 (mlsa.rs:54-79), next stage inputs
    const double* Xr = X + rx;
    const double2 y12 = *reinterpret_cast<const double2*>(Xr);
    const double2 y34 = *reinterpret_cast<const double2*>(Xr + 2);
    const double y5 = Xr[4];
    const double2 f12 = *reinterpret_cast<const double2*>(Xr + 8);
    const double2 f34 = *reinterpret_cast<const double2*>(Xr + 10);
    const double f5 = Xr[12];
    const double yprev = Xr[sm1], fprev = Xr[8 + sm1];
    const double e = In[g * 2 * kCH + inIdx];
    const double Af = __fma_rn(pp5, f5, __fma_rn(pp3, f34.x, pp1 * f12.x));
    const double Bf = __fma_rn(pp4, f34.y, pp2 * f12.y);
    const double Ay = __fma_rn(pp5, y5, __fma_rn(pp3, y34.x, pp1 * y12.x));
    const double By = __fma_rn(pp4, y34.y, pp2 * y12.y);
    const double d120 = (e + Af) - Bf;        // d12[0]
    const double x1 = __fma_rn(2.0, Af, e);   // x after df1
    const double d220 = (x1 + Ay) - By;       // d22[0]
    const double outv = __fma_rn(2.0, Ay, x1);
    gin = s == 0 ? d120 : fprev;
    uin = s == 0 ? d220 : yprev;
    if (head) Out[g * kCH + jout] = outv;
    inIdx = (inIdx + 1) & (2 * kCH - 1);
  };

  for (int i0 = 0; i0 < nsmp_max; i0 += kCH) {
    if (left == fp) {  // top of a frame (fp % kCH == 0): state taps for the cut check
      if (frames_done == my_own0 && my_cut) dump_state(0);
    }
#pragma unroll 1
    for (int j = 0; j < kCH; j += 2) {
      step(Cc0, C1, X0, j);
      step(Cc1, C0, X1, j + 1);
    }
    if (left == fp && frames_done == my_nfr) dump_state(1);
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
        if (n >= si.own0 * fp && n < si.nfr * fp) d.out[si.in_base + n] = Out[gg * kCH + i] * p.volume;
      }
      if ((chunk + 2) * kCH < nsmp_max) fill_chunk(chunk + 2);
    }
    __syncwarp();
  }
}

// Compares the two sides of every cut: state of segment g at the top of its own range (after its
// warm-up from zero) against the end state of the segment before it. max |difference| -> *worst.
__global__ void k_mlsa_seg_check(MlsaSegments sg, double* worst) {
  const int sid = blockIdx.x;
  const MlsaSegment s = sg.seg[sid];
  if (s.prev < 0) return;
  const double* A = sg.state + ((long long)sid * 2 + 0) * kMlsaSegStateDoubles;
  const double* B = sg.state + ((long long)s.prev * 2 + 1) * kMlsaSegStateDoubles;
  double m = 0.0;
  for (int i = threadIdx.x; i < 5 * 37; i += blockDim.x) {
    const double dd = fabs(A[i] - B[i]);
    m = (dd > m || dd != dd) ? (dd != dd ? __longlong_as_double(0x7ff0000000000000ll) : dd) : m;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0.0)
    atomicMax(reinterpret_cast<unsigned long long*>(worst), (unsigned long long)__double_as_longlong(m));
}

}  // namespace

int launch_mlsa_v3(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, const MlsaSegments& sg,
                   double* worst, cudaStream_t stream) {
  if (sg.n_seg == 0) return 0;
  if (p.nmcp != kNM || p.fperiod % kCH != 0 || d.out == nullptr || d.out == d.wave)
    return (int)cudaErrorInvalidValue;
  const int warps = (sg.n_seg + kSlots - 1) / kSlots;
  k_mlsa_v3<kWPB><<<(warps + kWPB - 1) / kWPB, 32 * kWPB, 0, stream>>>(b, p, d, sg);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  if (worst && sg.state) {
    e = cudaMemsetAsync(worst, 0, sizeof(double), stream);
    if (e != cudaSuccess) return (int)e;
    k_mlsa_seg_check<<<sg.n_seg, 64, 0, stream>>>(sg, worst);
    e = cudaGetLastError();
  }
  return (int)e;
}

}  // namespace jbk
