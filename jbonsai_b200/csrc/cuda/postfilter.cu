// Mel-cepstral post-filter (beta > 0), one warp per frame.
//
// Reference: MelCepstrum::postfilter_mcp (src/vocoder/cepstrum.rs:23-37), called once per frame by
// Vocoder::synthesize (src/vocoder/mod.rs:117-119) on the MLPG output before mc2b:
//     b  = mc2b(c);  e1 = b2en(b)
//     b[1] -= beta*alpha*b[2];  b[k] *= 1 + beta  (k >= 2);  e2 = b2en(b)
//     b[0] += ln(e1/e2) / 2;    c' = b2mc(b)          -> the vocoder then uses cc = mc2b(c')
// with b2en(b) = sum of squares of c2ir(freqt(b2mc(b), 575, -alpha), 576)
// (src/vocoder/coefficients.rs:65-78, cepstrum.rs:153-186). Per frame that is 2 x (35 x 576
// recurrence steps + 576^2/2 multiply-adds): ~1 MFLOP, four times the filter work of the frame
// itself, which is why it gets a kernel of its own instead of a line in k_frame_prep.
//
// The kernel rewrites VocoderDev::b in place: row f holds mc2b(c_f) on entry (k_frame_prep) and
// mc2b(postfilter_mcp(c_f)) on exit.
//
// freqt is a 35 x 576 grid recurrence, cell (i, j) needs (i, j-1), (i-1, j) and (i-1, j-1). Lane L
// owns rows 2L and 2L+1 and walks them along j, one column behind the lane before it (a wavefront
// of 576 + 36 steps, one 64-bit shuffle per step). Every cell is computed with the reference's
// operations in the reference's order, so the warped cepstrum is bit-identical to the oracle's.
//
// c2ir is ir[n] = (1/n) sum_{k=1..n} k c[k] ir[n-k]. Lane l owns six consecutive n (a block of 192
// per pass, three passes). For every already known ir[m] (shared-memory broadcast for m before the
// block, a shuffle from the owning lane inside the block) each lane updates its six sums; the six
// weights k c[k] it needs slide by one per m, so a step costs one shared-memory load for six FMAs.
// The sums run over m ascending with fused multiply-adds and a reciprocal multiply instead of the
// division: not the reference's rounding sequence, but the result only enters the waveform through
// ln(e1/e2)/2 (measured difference to the oracle's b[0]: ~1e-15).
#include "kernels.h"

namespace jbk {

namespace {

constexpr int kIr = 576;        // impulse response length of b2en (coefficients.rs:76)
constexpr int kNPL = 6;         // consecutive n per lane in c2ir
constexpr int kBlk = 32 * kNPL; // 192; kIr == 3 * kBlk
constexpr int kKcLen = kIr + kIr / kNPL + 8;
constexpr int kMaxM = 64;       // largest nmcp handled (two freqt rows per lane)
constexpr int kPfWarps = 8;

struct PfShared {
  double kc[kKcLen];  // k * g[k], stored at k + k/6 so that lanes 6 apart hit different banks
  double ir[kIr];
  double bs[kMaxM];   // coefficients b
  double mc[kMaxM];   // b2mc(b)
};

__device__ __forceinline__ int kpos(int i) { return i + i / kNPL; }

// b2en (coefficients.rs:75-78) of S.bs[0..M) by the whole warp.
__device__ double b2en_warp(PfShared& S, int M, double alpha, int lane) {
  const unsigned full = 0xffffffffu;
  // ---- b2mc (coefficients.rs:65-73)
  for (int i = lane; i < M; i += 32) S.mc[i] = i == M - 1 ? S.bs[i] : S.bs[i] + alpha * S.bs[i + 1];
  __syncwarp();

  // ---- freqt(m2 = 575, -alpha) (cepstrum.rs:153-173), wavefront over the rows
  const double af = -alpha, aa = 1.0 - af * af;
  const int rowA = 2 * lane, rowB = 2 * lane + 1;
  const double cA = rowA < M ? S.mc[rowA] : 0.0, cB = rowB < M ? S.mc[rowB] : 0.0;
  const int last_row = M - 1, last_lane = last_row >> 1;
  const bool last_is_b = (last_row & 1) != 0;
  double a1 = 0.0, a2 = 0.0;  // row A at columns jA-1, jA-2
  double b1 = 0.0;            // row B at column jB-1
  double in1 = 0.0;           // row above A at column jA-1
  double g0 = 0.0;
  const int steps = kIr + 2 * last_lane + (last_is_b ? 1 : 0);
  for (int s = 0; s < steps; ++s) {
    double x = __shfl_up_sync(full, b1, 1);  // row above A at column jA (computed one step ago)
    if (lane == 0) x = 0.0;                   // row -1: the zero-initialised cepstrum
    const int jA = s - 2 * lane, jB = jA - 1;
    double nA = 0.0, nB = 0.0;
    if (jA >= 0) nA = jA == 0 ? cA + af * x : (jA == 1 ? aa * in1 + af * x : in1 + af * (x - a1));
    if (jB >= 0) nB = jB == 0 ? cB + af * a1 : (jB == 1 ? aa * a2 + af * a1 : a2 + af * (a1 - b1));
    if (lane == last_lane) {
      const int j = last_is_b ? jB : jA;
      const double v = last_is_b ? nB : nA;
      if (j >= 0 && j < kIr) {
        S.kc[kpos(j)] = (double)j * v;  // k as f64 * self[k] (cepstrum.rs:181); kc[0] = 0
        if (j == 0) g0 = v;
      }
    }
    a2 = a1; a1 = nA; b1 = nB; in1 = x;
  }
  g0 = __shfl_sync(full, g0, last_lane);
  __syncwarp();

  // ---- c2ir(576) (cepstrum.rs:175-186) and the energy
  const double ir0 = exp(g0);
  double e = 0.0;
  auto kcz = [&](int i) -> double { return S.kc[kpos(i > 0 ? i : 0)]; };
  for (int nb = 0; nb < kIr; nb += kBlk) {
    const int idx0 = nb + kNPL * lane;  // n of this lane's first sum
    double d[kNPL], w[kNPL], rc[kNPL];
#pragma unroll
    for (int r = 0; r < kNPL; ++r) {
      d[r] = 0.0;
      w[r] = kcz(idx0 + r);
      rc[r] = 1.0 / (double)(idx0 + r);
    }
    // ir[m] known before this block: m = 0 .. nb-1
    for (int m0 = 0; m0 < nb; m0 += kNPL) {
#pragma unroll
      for (int u = 0; u < kNPL; ++u) {
        const double v = S.ir[m0 + u];
#pragma unroll
        for (int r = 0; r < kNPL; ++r) d[r] = __fma_rn(w[(r - u + kNPL) % kNPL], v, d[r]);
        w[(kNPL - 1 - u) % kNPL] = kcz(idx0 - (m0 + u) - 1);
      }
    }
    // inside the block: lane q finishes its six values one after the other
    for (int q = 0; q < 32; ++q) {
#pragma unroll
      for (int u = 0; u < kNPL; ++u) {
        const int n = nb + kNPL * q + u;
        const double val = n == 0 ? ir0 : d[u] * rc[u];
        const double v = __shfl_sync(full, val, q);
        if (lane == q) {
          S.ir[n] = v;
          e = __fma_rn(v, v, e);
        }
#pragma unroll
        for (int r = 0; r < kNPL; ++r) d[r] = __fma_rn(w[(r - u + kNPL) % kNPL], v, d[r]);
        w[(kNPL - 1 - u) % kNPL] = kcz(idx0 - n - 1);
      }
    }
    __syncwarp();
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(full, e, o);
  return e;
}

__global__ void __launch_bounds__(32 * kPfWarps)
k_postfilter(VocoderParams p, VocoderDev d, double beta, int64_t total_frames) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  PfShared& S = reinterpret_cast<PfShared*>(pf_smem)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const int64_t gf = (int64_t)blockIdx.x * kPfWarps + (threadIdx.x >> 5);
  if (gf >= total_frames) return;  // whole warp; warps of a block never sync with each other
  const int M = p.nmcp;
  const double alpha = p.alpha;
  double* row = d.b + gf * M;
  for (int i = lane; i < M; i += 32) S.bs[i] = row[i];
  __syncwarp();

  const double e1 = b2en_warp(S, M, alpha, lane);
  if (lane == 0) S.bs[1] -= beta * alpha * S.bs[2];
  for (int k = 2 + lane; k < M; k += 32) S.bs[k] *= 1.0 + beta;
  __syncwarp();
  const double e2 = b2en_warp(S, M, alpha, lane);
  if (lane == 0) S.bs[0] += log(e1 / e2) / 2.0;
  __syncwarp();
  // c' = b2mc(b) (cepstrum.rs:35), then cc = mc2b(c') as mod.rs:119 does
  for (int i = lane; i < M; i += 32) S.mc[i] = i == M - 1 ? S.bs[i] : S.bs[i] + alpha * S.bs[i + 1];
  __syncwarp();
  if (lane == 0) {
    double nxt = S.mc[M - 1];
    row[M - 1] = nxt;
    for (int i = M - 2; i >= 0; --i) {
      nxt = (alpha != 0.0) ? S.mc[i] - alpha * nxt : S.mc[i];
      row[i] = nxt;
    }
  }
}

// b_first[u] = row of the first frame of utterance u (before the post-filter rewrites it)
__global__ void k_first_rows(BatchLayout b, int M, const double* rows, double* b_first) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= b.n_utt * M) return;
  const int u = i / M, k = i - u * M;
  if (b.frame_off[u + 1] > b.frame_off[u]) b_first[i] = rows[b.frame_off[u] * M + k];
}

}  // namespace

int launch_postfilter(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, double beta,
                      double* b_first, int64_t total_frames, cudaStream_t stream) {
  if (!(beta > 0.0) || p.nmcp <= 2 || total_frames == 0) return 0;  // cepstrum.rs:24
  if (p.nmcp > kMaxM || b_first == nullptr) return (int)cudaErrorInvalidValue;
  const int n = b.n_utt * p.nmcp;
  k_first_rows<<<(n + 255) / 256, 256, 0, stream>>>(b, p.nmcp, d.b, b_first);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  static_assert(kIr == 3 * kBlk, "c2ir passes");
  const size_t smem = sizeof(PfShared) * kPfWarps;
  e = cudaFuncSetAttribute(k_postfilter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  const int64_t blocks = (total_frames + kPfWarps - 1) / kPfWarps;
  k_postfilter<<<(unsigned)blocks, 32 * kPfWarps, smem, stream>>>(p, d, beta, total_frames);
  return (int)cudaGetLastError();
}

}  // namespace jbk
