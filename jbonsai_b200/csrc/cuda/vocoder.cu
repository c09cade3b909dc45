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

#include <cstdlib>

#include "kernels.h"
#include "cnt_advance.h"
#include "cr_exp.h"

namespace jbk {

namespace {

#include "exp_table.inc"

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
// what: bit 0 = pitch period from lf0, bit 1 = b = mc2b(mcp) (the two halves depend on different streams)
__global__ void k_frame_prep(VocoderParams p, VocoderDev d, int64_t total_frames, int what) {
  int64_t gf = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gf >= total_frames) return;
  if (what & 1) {
    const double lf0 = d.lf0[gf];
    double pitch = 0.0;
    if (lf0 != kNoData) {
      double l = lf0 < kMinLf0 ? kMinLf0 : (lf0 > kMaxLf0 ? kMaxLf0 : lf0);
      // exp rounded correctly (cr_exp.h): the period, and with it every pulse position, is then the one the host
      // libm gives wherever that rounds correctly too (99.93 % of the arguments; tools/exp_sweep.py)
      pitch = p.rate / cr_exp(l, c_exp2_64, kLn2_64_hi, kLn2_64_mid, kLn2_64_lo, k64_Ln2);
    }
    d.pitch[gf] = pitch;
  }
  if (!(what & 2)) return;
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

// ---------------------------------------------------------------- frame prep + increments, fused
// b = mc2b(mcp) (cepstrum.rs:139-149) and binc[f] = (b[f] - b[f-1]) / fperiod (vocoder/mod.rs:118-127) in one
// pass over the spectrum: k_frame_prep (a thread walks its own 280-byte row: 1 TB/s) + k_row_increments (reads b
// back, one binary search per element) cost 0.57 + 0.29 ms on the bench batch. Here a block stages the rows of
// kCoefFrames frames and of the frame before them in shared memory with coalesced loads, a thread runs the mc2b
// recurrence on a row (odd row stride: no bank conflicts), and rows and increments leave coalesced. Same
// operations per element as the two kernels, so the same bits. Stage 0 without post-filter only (no b_first: the
// first frame of an utterance starts from its own row, i.e. a zero increment).
constexpr int kCoefFrames = 128;
__global__ void __launch_bounds__(kCoefFrames) k_coef_rows(BatchLayout b, VocoderParams p, VocoderDev d,
                                                           int64_t total_frames) {
  extern __shared__ double tile[];  // [kCoefFrames + 1][M | 1]: row 0 = the frame before the block's first
  __shared__ unsigned char first[kCoefFrames];
  const int M = p.nmcp, RS = M | 1;
  const int64_t f0 = (int64_t)blockIdx.x * kCoefFrames;
  const int nf = (int)min((int64_t)kCoefFrames, total_frames - f0);
  const int tid = threadIdx.x;
  const int64_t base = (f0 > 0 ? f0 - 1 : 0) * M;  // first element staged
  const int r0 = f0 > 0 ? 0 : 1;                   // its row in the tile
  const int n_in = (nf + 1 - r0) * M;
  for (int e = tid; e < n_in; e += kCoefFrames) {
    const int row = e / M;
    tile[(row + r0) * RS + (e - row * M)] = d.mcp[base + e];
  }
  if (tid < nf) {
    const int64_t gf = f0 + tid;
    first[tid] = gf == b.frame_off[find_utt(b.frame_off, b.n_utt, gf)];
  }
  __syncthreads();
  for (int row = r0 + tid; row <= nf; row += kCoefFrames) {  // nf + 1 rows over 128 threads: thread 0 takes two
    double* c = tile + row * RS;
    double nxt = c[M - 1];
    for (int i = M - 2; i >= 0; --i) {
      nxt = (p.alpha != 0.0) ? c[i] - p.alpha * nxt : c[i];
      c[i] = nxt;
    }
  }
  __syncthreads();
  const double fp = (double)p.fperiod;
  for (int e = tid; e < nf * M; e += kCoefFrames) {
    const int fr = e / M, k = e - fr * M;
    const double cur = tile[(fr + 1) * RS + k];
    const double prev = first[fr] ? cur : tile[fr * RS + k];
    d.b[f0 * M + e] = cur;
    d.binc[f0 * M + e] = (cur - prev) / fp;
  }
}

// ---------------------------------------------------------------- pulse train
// One thread per utterance walks the pitch counter sample by sample (the additions must round
// exactly like excitation.rs:72-86, so the walk is sequential) but only WRITES the pulses: the
// pulse array is zeroed beforehand (launch_pulse memsets it; dense=1 zeroes frame by frame for the
// chunked streaming path). Pitch values of the next frame are fetched one frame ahead.
__global__ void __launch_bounds__(32) k_pulse(BatchLayout b, VocoderParams p, VocoderDev d, int dense) {
  int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= b.n_utt) return;
  if (b.order) u = b.order[u];  // neighbouring lanes get utterances of similar length
  const int64_t foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  int f0 = d.frame_begin, f1 = f0 + d.frame_count;
  if (f1 > T || f1 < f0) f1 = T;
  double p_cur = 0.0, cnt = 0.0;
  if (d.state && f0 > 0) {
    p_cur = d.state[(int64_t)u * kVocStateDoubles + 223];
    cnt = d.state[(int64_t)u * kVocStateDoubles + 224];
  }
  const int fp = p.fperiod;
  const double fper = (double)fp;
  double pitch_next = f0 < f1 ? d.pitch[foff + f0] : 0.0;
  for (int f = f0; f < f1; ++f) {
    const double pitch = pitch_next;
    if (f + 1 < f1) pitch_next = d.pitch[foff + f + 1];
    double inc;
    if (p_cur != 0.0 && pitch != 0.0) {  // Excitation::start (excitation.rs:25-33)
      inc = (pitch - p_cur) / fper;
    } else {
      inc = 0.0;
      p_cur = pitch;
      cnt = pitch;
    }
    double* out = d.pulse + (foff + f) * fp;
    if (dense)
      for (int i = 0; i < fp; ++i) out[i] = 0.0;
    if (p_cur != 0.0) {
      for (int i = 0; i < fp; ++i) {  // voiced_frame (excitation.rs:72-86)
        cnt += 1.0;
        if (cnt >= p_cur) {
          cnt -= p_cur;
          out[i] = sqrt(p_cur);
        }
        p_cur += inc;
      }
    }
    p_cur = pitch;  // Excitation::end (excitation.rs:102-104)
  }
  if (d.state) {
    d.state[(int64_t)u * kVocStateDoubles + 223] = p_cur;
    d.state[(int64_t)u * kVocStateDoubles + 224] = cnt;
  }
}

// Whole-utterance synthesis: the pitch counter restarts at every unvoiced -> voiced transition
// (Excitation::start resets pitch_of_curr_point and pitch_counter when the previous frame was
// unvoiced, excitation.rs:25-33), so each maximal run of voiced frames is an independent walk.
// k_pulse_list collects the first frame of every run (d.runs[0] = count, d.runs[1..] = frames, in
// no particular order); k_pulse_runs walks one run per thread, so that the 32 lanes of a warp all
// carry a walk (with one thread per frame a warp held one walking lane and 31 idle ones, and the
// kernel was bound by the issue slots of ~6000 such warps).
__global__ void __launch_bounds__(256) k_pulse_list(BatchLayout b, VocoderDev d, int64_t total_frames) {
  const int64_t gf = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gf >= total_frames) return;
  if (d.pitch[gf] == 0.0) return;
  if (gf > 0 && d.pitch[gf - 1] != 0.0 && b.frame_off[find_utt(b.frame_off, b.n_utt, gf)] != gf) return;
  const int slot = atomicAdd(d.runs, 1);
  d.runs[1 + slot] = (int32_t)gf;
}

// Walks per warp. Measured inside the MLPG phase (the walks share the chip with k_mlpg_gv of the MCP stream):
// 4 -> 6.29 ms, 8 -> 6.11 ms, 16 -> 6.45 ms for the phase; ncu: with 4, 1.66 threads are active per warp
// instruction on average and the kernel issues 458 M warp instructions.
#ifndef JB_RUN_LANES
#define JB_RUN_LANES 8
#endif
constexpr int kRunLanes = JB_RUN_LANES;
__global__ void __launch_bounds__(128) k_pulse_runs(BatchLayout b, VocoderParams p, VocoderDev d,
                                                    int64_t total_frames) {
  // kRunLanes walks per warp: a walk leaves the 8-sample fast path about once in 30 blocks, and a
  // warp pays for the slow path whenever any of its lanes does, so fewer walks per warp (and more
  // warps) finish sooner than 32 walks in lockstep.
  if (threadIdx.x % (32 / kRunLanes) != 0) return;
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / (32 / kRunLanes);
  if (i >= d.runs[0]) return;
  const int64_t gf = d.runs[1 + i];
  const double first = d.pitch[gf];
  const int u = find_utt(b.frame_off, b.n_utt, gf);
  const int64_t foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  const int f = (int)(gf - foff);
  const int fp = p.fperiod;
  const double fper = (double)fp;
  double p_cur = first, cnt = first, inc = 0.0, pitch = first;
  for (int ff = f;;) {
    double* out = d.pulse + (foff + ff) * fp;
    const double pitch_next = ff + 1 < T ? d.pitch[foff + ff + 1] : 0.0;
    int i = 0;
    // Eight samples at a time. The pitch chain p += inc is independent of the counter; the counter
    // itself moves by exact integer steps while it stays inside one binade (cnt >= 1, same exponent
    // for cnt and cnt + 8): then cnt + k is bit for bit the k-fold "+= 1.0" of the reference, and a
    // block in which no sample reaches its period is just cnt += 8. Any other block (a pulse, or a
    // binade crossing where "+= 1.0" rounds) is replayed sample by sample below.
    for (; i + 8 <= fp; i += 8) {
      double pk[8];
      pk[0] = p_cur;
#pragma unroll
      for (int k = 1; k < 8; ++k) pk[k] = pk[k - 1] + inc;
      const double c8 = cnt + 8.0;
      bool slow = !(cnt >= 1.0) || (__double2hiint(c8) >> 20) != (__double2hiint(cnt) >> 20);
#pragma unroll
      for (int k = 0; k < 8; ++k) slow |= (cnt + (double)(k + 1)) >= pk[k];
      if (!slow) {
        cnt = c8;
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) {  // voiced_frame (excitation.rs:72-86)
          cnt += 1.0;
          if (cnt >= pk[k]) {
            cnt -= pk[k];
            out[i + k] = sqrt(pk[k]);
          }
        }
      }
      p_cur = pk[7] + inc;
    }
    for (; i < fp; ++i) {  // voiced_frame (excitation.rs:72-86)
      cnt += 1.0;
      if (cnt >= p_cur) {
        cnt -= p_cur;
        out[i] = sqrt(p_cur);
      }
      p_cur += inc;
    }
    p_cur = pitch;  // Excitation::end
    if (pitch_next == 0.0) break;
    ++ff;
    pitch = pitch_next;
    inc = (pitch - p_cur) / fper;  // Excitation::start with both pitches non-zero
  }
}

// ---------------------------------------------------------------- pulse walks, one warp per run
// k_pulse_runs gives a run to one lane, and that lane pays, per sample, a dependent addition for the pitch chain and
// one for the counter: ~3500 cycles per frame, so the longest voiced run of the batch (hundreds of frames) sets the
// kernel's time (2.7 ms on the bench batch with 6 % of the warps active). Two things are not serial, though:
//  * the pitch chain restarts at every frame (Excitation::end snaps p to the frame's pitch), so the 32 lanes run
//    the chains of 32 frames side by side and keep every eighth value in shared memory;
//  * between two pulses the counter moves by exact steps: "+= 1.0" is exact while the sum stays in the binade of the
//    counter and rounds only on the step that crosses into the next one (cnt_advance replays exactly those steps),
//    so the counter at the start of any 8-sample block is known WITHOUT walking, as long as no pulse lies before it.
// So a warp tests 32 blocks (256 samples) at once, each lane replaying its own block sample by sample with the
// reference's operations (excitation.rs:72-86); the first lane that sees a pulse is right (nothing precedes it), it
// writes its pulses and hands its counter to the lanes behind it, which start over from there. About two rounds
// per 256 samples instead of 256 dependent steps; same bits.
// cnt_advance (cnt_advance.h): the counter after m steps, without walking; host-testable (tests/test_cnt_advance.py).

constexpr int kPwWarps = 4;
constexpr int kPwMaxBlocks = 32;  // 8-sample blocks per frame: fperiod <= 256
struct PulseWarpShared {
  double pk8[32][kPwMaxBlocks + 1];  // [frame of the batch][block]: the pitch chain at the block's first sample
  double inc[32];
};

// Persistent grid: a warp fetches the next run from the counter `next` (zeroed by the launcher) until the list is done.
__global__ void __launch_bounds__(32 * kPwWarps) k_pulse_warp(BatchLayout b, VocoderParams p, VocoderDev d, int* next) {
  __shared__ PulseWarpShared sh[kPwWarps];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  PulseWarpShared& S = sh[wib];
  const unsigned full = 0xffffffffu;
  const int n_runs = d.runs[0];
 for (;;) {
  int run = 0;
  if (lane == 0) run = atomicAdd(next, 1);
  run = __shfl_sync(full, run, 0);
  if (run >= n_runs) return;
  __syncwarp();  // the tables of the run before are free
  const int64_t gf0 = d.runs[1 + run];
  const int u = find_utt(b.frame_off, b.n_utt, gf0);
  const int64_t foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  const int f0 = (int)(gf0 - foff);
  const int fp = p.fperiod, NB = fp >> 3;
  const double fper = (double)fp;
  const double* const pitch = d.pitch + foff;
  double cnt = pitch[f0];  // Excitation::start from an unvoiced frame: the counter starts at the period
  for (int fb = f0;; fb += 32) {
    // ---- frames fb .. fb + nvalid - 1 of this batch belong to the run
    const int f = fb + lane;
    const double pf = f < T ? pitch[f] : 0.0;
    const unsigned unv = __ballot_sync(full, pf == 0.0);
    const int nvalid = unv ? __ffs(unv) - 1 : 32;
    if (nvalid == 0) break;
    if (lane < nvalid) {
      const bool first = f == f0;
      double pc = first ? pf : pitch[f - 1];  // Excitation::end left p at the pitch of the frame before
      const double inc = first ? 0.0 : (pf - pc) / fper;
      S.inc[lane] = inc;
      for (int i = 0; i < NB; ++i) {
        S.pk8[lane][i] = pc;
#pragma unroll
        for (int k = 0; k < 8; ++k) pc += inc;
      }
    }
    __syncwarp();
    // ---- 32 blocks per step
    const int nblk = nvalid * NB;
    for (int g = 0; g < nblk; g += 32) {
      const int blk = g + lane;
      const bool act = blk < nblk;
      const int fl = act ? blk / NB : 0, bi = act ? blk - fl * NB : 0;
      const double pk0 = S.pk8[fl][bi], inc = S.inc[fl];
      const int nact = min(32, nblk - g);
      int base = 0;
      const double lb = (inc < 0.0 ? pk0 + 8.0 * inc : pk0) - 1e-3;  // no sample of the block has a period below this
      while (base < nact) {
        // Cheap test first: the counter at the end of this block is cnt + 8 (lane - base + 1) up to a few ulps of
        // rounding; a block that stays a safe distance below its smallest period holds no pulse, and its lane
        // needs no exact counter. Only the other lanes (one or two per round) replay their block.
        const bool mine = act && lane >= base && !(cnt + (double)(8 * (lane - base + 1)) < lb);
        double c = cnt, pk[8];
        unsigned mask = 0;
        if (mine) {
          c = cnt_advance(cnt, 8 * (lane - base));
          pk[0] = pk0;
#pragma unroll
          for (int k = 1; k < 8; ++k) pk[k] = pk[k - 1] + inc;
#pragma unroll
          for (int k = 0; k < 8; ++k) {  // voiced_frame (excitation.rs:72-86)
            c += 1.0;
            if (c >= pk[k]) {
              c -= pk[k];
              mask |= 1u << k;
            }
          }
        }
        const unsigned hit = __ballot_sync(full, mask != 0);
        if (hit == 0) {
          cnt = cnt_advance(cnt, 8 * (nact - base));  // no pulse in the rest of the step
          break;
        }
        const int L = __ffs(hit) - 1;
        if (lane == L) {
          double* out = d.pulse + ((foff + fb + fl) * fp + bi * 8);
#pragma unroll
          for (int k = 0; k < 8; ++k)
            if (mask >> k & 1u) out[k] = sqrt(pk[k]);
        }
        cnt = __shfl_sync(full, c, L);
        base = L + 1;
      }
    }
    if (nvalid < 32) break;
    __syncwarp();  // the tables are rewritten for the next batch of frames
  }
 }
}

// ---------------------------------------------------------------- excitation + gain
// One block per frame. Gather form of the ring-buffer overlap-add (excitation.rs:43-64,67-100):
// output sample n sums, oldest source first, over source samples m = n-(nlpf-1) .. n:
//    [m == n-h]   + noise[m]                       (get_antipode_mut, h = (nlpf-1)/2)
//    [voiced(m)]  + (pulse[m]-noise[m]) * lpf_{frame(m)}[n-m]
// which is exactly the order in which the reference's scatter adds reach ring slot n. Unvoiced
// sources contribute (pulse-noise) := 0, i.e. an exact "+ 0.0". The gain exp(c[0]) (mod.rs:129-131)
// uses coefficients[0] advanced by repeated addition like the per-sample loop (mod.rs:133-135);
// that 240-step chain runs in the last thread of the block while the others do the taps.
constexpr int kMaxLpf = 64;
constexpr int kMaxFperiod = 1024;

__global__ void __launch_bounds__(256) k_excitation(BatchLayout b, VocoderParams p, VocoderDev d,
                                                    int64_t total_frames) {
  const int64_t gf = blockIdx.x;
  if (gf >= total_frames) return;
  const int u = find_utt(b.frame_off, b.n_utt, gf);
  const int64_t foff = b.frame_off[u];
  const int f = (int)(gf - foff);
  if (f < d.frame_begin || (d.frame_count >= 0 && f - d.frame_begin >= d.frame_count)) return;

  __shared__ double s_diff[kMaxFperiod + kMaxLpf];   // voiced(m) ? pulse[m] - noise[m] : 0
  __shared__ double s_noise[kMaxFperiod + kMaxLpf];
  __shared__ double s_lpf[2 * kMaxLpf];               // [0..64) previous frame, [64..128) this frame
  __shared__ double s_c0[kMaxFperiod];
  const int nl = p.nlpf, fp = p.fperiod, hist = nl > 0 ? nl - 1 : 0;
  const int64_t n0 = (int64_t)f * fp;  // first sample of this frame within the utterance
  const bool voiced_cur = d.pitch[gf] != 0.0;
  const bool voiced_prev = f > 0 && d.pitch[gf - 1] != 0.0;
  const double* pulse = d.pulse + foff * fp + n0 - hist;
  const double* noise = d.noise + n0 - hist;
  if (nl == 0 && !voiced_cur) {
    // no LPF stream (excitation.rs:87-99): noise is drawn on unvoiced samples only, so this frame
    // reads the sequence at fperiod * (unvoiced frames before it)
    int before = 0;
    for (int j0 = 0; j0 < f; j0 += blockDim.x) {
      const int j = j0 + threadIdx.x;
      before += __syncthreads_count(j < f && d.pitch[foff + j] == 0.0);
    }
    noise = d.noise + (int64_t)before * fp;
  }

  for (int i = threadIdx.x; i < fp + hist; i += blockDim.x) {
    const bool in = n0 - hist + i >= 0;
    const double nz = in ? noise[i] : 0.0;
    const bool v = i >= hist ? voiced_cur : voiced_prev;
    s_noise[i] = nz;
    s_diff[i] = (in && v) ? (nl > 0 ? pulse[i] - nz : pulse[i]) : 0.0;
  }
  for (int i = threadIdx.x; i < nl; i += blockDim.x) {
    s_lpf[kMaxLpf + i] = d.lpf[gf * nl + i];
    s_lpf[i] = f > 0 ? d.lpf[(gf - 1) * nl + i] : 0.0;
  }
  if (threadIdx.x == blockDim.x - 1) {
    const double cc = d.b[gf * p.nmcp];
    double c = f > 0 ? d.b[(gf - 1) * p.nmcp] : (d.b_first ? d.b_first[(int64_t)u * p.nmcp] : cc);
    const double cinc = (cc - c) / (double)fp;
    for (int i = 0; i < fp; ++i) {
      s_c0[i] = c;
      c += cinc;
    }
  }
  __syncthreads();

  double* wave = d.wave + foff * fp + n0;
  for (int i = threadIdx.x; i < fp; i += blockDim.x) {
    double x;
    if (nl > 0) {
      const int h = (nl - 1) / 2;
      if (!voiced_cur && !(voiced_prev && i < hist)) {
        x = 0.0 + s_noise[i + hist - h];  // only the centre-tap noise reaches this slot
      } else {
        x = 0.0;
        // sources of the previous frame (k > i) use its filter, the rest this frame's
        for (int k = nl - 1; k >= 0; --k) {
          const int j = i + hist - k;
          if (k == h) x += s_noise[j];
          x += s_diff[j] * s_lpf[(k <= i ? kMaxLpf : 0) + k];
        }
      }
    } else {
      x = voiced_cur ? s_diff[i] : s_noise[i];
    }
    if (p.stage != 0) x *= s_c0[i];  // Stage::NonZero gain (mod.rs:163)
    else if (x != 0.0) x *= exp(s_c0[i]);
    wave[i] = x;
  }
}

// Fast path of the same computation for nlpf = NL <= 32 and 8 | fperiod <= 256 (every shipped
// voice: 31 taps, 240 samples): one warp per frame, each lane owns eight consecutive samples and
// slides a 38-value register window over them, so a multiply-add costs no shared-memory read of its
// own (the block-per-frame kernel above reads two operands per tap and is shared-memory bound).
// Same operands, same order of additions per sample. The staging arrays skip one slot in nine so
// that lanes reading with a stride of eight doubles fall into different banks.
#ifndef JB_EX_WARPS
#define JB_EX_WARPS 4
#endif
constexpr int kExWarps = JB_EX_WARPS;
// Frames one warp handles in a row. coefficients[0] of every sample is a chain of fperiod dependent additions
// (mod.rs:133-135: the rounding sequence of the reference); run by one lane per frame it cost 240 warp-wide
// DADDs and 2000 cycles of latency for every frame. Here lane q runs the chain of frame q of the group and
// keeps every eighth value; the lane that owns samples 8 L .. 8 L + 7 of a frame restarts from the exact value
// at 8 L: 240 / G + 7 additions per frame instead of 240, same bits. Measured (B200, bench batch): one frame
// per warp 3.24 ms; G = 16: 2.40, 8: 2.09, 4: 2.05 ms (a larger group costs shared memory, i.e. resident warps).
#ifndef JB_EX_GROUP
#define JB_EX_GROUP 8
#endif
constexpr int kExGroup = JB_EX_GROUP;
template <int NL>
__global__ void __launch_bounds__(32 * kExWarps) k_excitation_w(BatchLayout b, VocoderParams p, VocoderDev d,
                                                                int64_t total_frames) {
  constexpr int HIST = NL - 1, H = (NL - 1) / 2, R = 8, WIN = R + HIST;
  constexpr int MAXN = 256 + HIST, PADN = MAXN + MAXN / 8 + 2;
  __shared__ double s_diff[kExWarps][PADN];
  __shared__ double s_noise[kExWarps][PADN];
  __shared__ double s_lpf[kExWarps][2][32];  // [0] previous frame's filter, [1] this frame's
  __shared__ double s_c8[kExWarps][kExGroup][33];  // coefficients[0] at samples 0, 8, 16, ... of every frame
  __shared__ double s_cinc[kExWarps][kExGroup];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Work items: every frame of the batch, or (chunked synthesis: d.frame_count >= 0) only frames
  // [frame_begin, frame_begin + frame_count) of every utterance, numbered utterance-major.
  const bool chunked = d.frame_count >= 0;
  const int64_t n_items = chunked ? (int64_t)b.n_utt * d.frame_count : total_frames;
  const int64_t v0 = ((int64_t)blockIdx.x * kExWarps + w) * kExGroup;
  if (v0 >= n_items) return;  // whole warp; warps never sync with each other
  const int fp = p.fperiod;
  auto pad = [](int j) { return j + (j >> 3); };
  auto locate = [&](int64_t v, int& u, int& f, int64_t& gf) -> bool {
    if (v >= n_items) return false;
    if (chunked) {
      u = (int)(v / d.frame_count);
      f = d.frame_begin + (int)(v - (int64_t)u * d.frame_count);
      gf = b.frame_off[u] + f;
      return gf < b.frame_off[u + 1];
    }
    gf = v;
    u = find_utt(b.frame_off, b.n_utt, gf);
    f = (int)(gf - b.frame_off[u]);
    return f >= d.frame_begin;
  };

  if (lane < kExGroup) {  // the chain of item v0 + lane (mod.rs:133-135)
    int u, f;
    int64_t gf;
    if (locate(v0 + lane, u, f, gf)) {
      const double cc = d.b[gf * p.nmcp];
      double c = f > 0 ? d.b[(gf - 1) * p.nmcp] : (d.b_first ? d.b_first[(int64_t)u * p.nmcp] : cc);
      const double cinc = (cc - c) / (double)fp;
      s_cinc[w][lane] = cinc;
      for (int i = 0; i < fp; i += R) {
        s_c8[w][lane][i / R] = c;
#pragma unroll
        for (int r = 0; r < R; ++r) c += cinc;
      }
    }
  }
  __syncwarp();

  for (int q = 0; q < kExGroup; ++q) {
    if (v0 + q >= n_items) break;  // uniform
    int u, f;
    int64_t gf;
    if (!locate(v0 + q, u, f, gf)) continue;  // uniform
    const int64_t foff = b.frame_off[u];
    const int64_t n0 = (int64_t)f * fp;
    const bool voiced_cur = d.pitch[gf] != 0.0;
    const bool voiced_prev = f > 0 && d.pitch[gf - 1] != 0.0;
    const double* pulse = d.pulse + foff * fp + n0 - HIST;
    const double* noise = d.noise + n0 - HIST;

    __syncwarp();  // the previous frame's staging arrays have been read
    for (int i = lane; i < fp + HIST; i += 32) {
      const bool in = n0 - HIST + i >= 0;
      const double nz = in ? noise[i] : 0.0;
      const bool v = i >= HIST ? voiced_cur : voiced_prev;
      s_noise[w][pad(i)] = nz;
      s_diff[w][pad(i)] = (in && v) ? pulse[i] - nz : 0.0;
    }
    if (lane < NL) {
      s_lpf[w][1][lane] = d.lpf[gf * NL + lane];
      s_lpf[w][0][lane] = f > 0 ? d.lpf[(gf - 1) * NL + lane] : 0.0;
    }
    __syncwarp();

    const int i0 = lane * R;
    if (i0 < fp) {
      double* wave = d.wave + foff * fp + n0 + i0;
      double x[R];
      if (!voiced_cur && !(voiced_prev && i0 < HIST)) {
#pragma unroll
        for (int r = 0; r < R; ++r) x[r] = 0.0 + s_noise[w][pad(i0 + r + HIST - H)];  // centre-tap noise only
      } else {
        double win[WIN], nz[R];
#pragma unroll
        for (int t = 0; t < WIN; ++t) win[t] = s_diff[w][pad(i0 + t)];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          nz[r] = s_noise[w][pad(i0 + r + HIST - H)];
          x[r] = 0.0;
        }
#pragma unroll
        for (int k = NL - 1; k >= 0; --k) {
          const double lc = s_lpf[w][1][k], lp = s_lpf[w][0][k];
#pragma unroll
          for (int r = 0; r < R; ++r) {
            if (k == H) x[r] += nz[r];
            // sources of the previous frame (k > i) use its filter, the rest this frame's
            x[r] += win[r + HIST - k] * ((k - r > i0) ? lp : lc);
          }
        }
      }
      double c0 = s_c8[w][q][lane];  // coefficients[0] at sample i0, then by repeated addition as the chain
      const double cinc = s_cinc[w][q];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        double v = x[r];
        if (p.stage != 0) v *= c0;  // Stage::NonZero gain (mod.rs:163)
        else if (v != 0.0) v *= exp(c0);  // mod.rs:129-131
        wave[r] = v;
        c0 += cinc;
      }
    }
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
      // is_first: c = mc2b(first frame), before the post-filter (mod.rs:79-88)
      const double c = !factive ? cc : (f == 0 ? (d.b_first ? d.b_first[(int64_t)uu * NM + k] : cc) : s_c[k][sg]);
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

// ---------------------------------------------------------------- MLSA, any vector length
// MelLogSpectrumApproximation::new(nmcp) and fir are generic in the vector length (mlsa.rs:38-45,128-163); the
// fast kernels keep the delay lines in registers and exist for the lengths they were instantiated for. This one
// takes any 2 <= nmcp <= kAnyMaxNM and any fperiod: one warp per utterance, lane i - 1 owns Pade stage i, its
// delay line lives in shared memory and the taps run one at a time in the form of the reference's tail loop
// (mlsa.rs:155-161; the four-tap blocks of :134-153 are the same recurrence unrolled). Stage outputs are combined
// through shuffles in the reference's order (i = 5..1) as in variant 0. Slow (one serial chain per utterance) but
// never JB_ERR_UNSUPPORTED. Streaming state (nmcp <= 35: the layout of kernels.h with NM - 1 taps per stage).
constexpr int kAnyMaxNM = 128;
__device__ __forceinline__ double fir_smem(double* dl, int NM, double x, double a, double iaa, const double* c) {
  double rem = -a * x + dl[0];
  dl[0] = iaa * x + a * dl[0];
  double y = 0.0;
  for (int k = 2; k < NM; ++k) {
    const double dv = dl[k - 1];
    const double nd = iaa * rem + a * dv;
    rem = -a * rem + dv;
    dl[k - 1] = nd;
    y += c[k] * nd;
  }
  return y;
}

__global__ void __launch_bounds__(32) k_mlsa_any(BatchLayout b, VocoderParams p, VocoderDev d) {
  extern __shared__ double any_sm[];
  const int NM = p.nmcp, NT = NM - 1;
  double* const s_dl = any_sm;           // [5][NT]
  double* const s_c = s_dl + 5 * NT;     // [NM] interpolated coefficients
  double* const s_cinc = s_c + NM;
  double* const s_cc = s_cinc + NM;
  const int lane = threadIdx.x;
  const int u = b.order ? b.order[blockIdx.x] : (int)blockIdx.x;
  const int64_t foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  int f0 = d.frame_begin, f1 = f0 + d.frame_count;
  if (f1 > T || f1 < f0) f1 = T;
  if (f0 >= f1) return;
  const int st = lane + 1;
  const bool stage_lane = lane < 5;
  const double a = p.alpha, iaa = 1.0 - a * a;
  const double pp = stage_lane ? c_ppade[st] : 0.0;
  double* const dl = s_dl + (stage_lane ? lane : 0) * NT;
  for (int k = lane; k < 5 * NT; k += 32) s_dl[k] = 0.0;
  double d11 = 0.0, d12 = 0.0, d22 = 0.0, in1 = 0.0, in2 = 0.0;
  const int T5 = 5 * NT;  // state layout: taps, then d11[6], d12[6], d22[6], coefficients[NM]
  double* stp = d.state ? d.state + (int64_t)u * kVocStateDoubles : nullptr;
  if (stp != nullptr && f0 > 0) {
    for (int k = lane; k < T5; k += 32) s_dl[k] = stp[k];
    for (int k = lane; k < NM; k += 32) s_c[k] = stp[T5 + 18 + k];
    if (stage_lane) {
      d11 = stp[T5 + st];
      d12 = stp[T5 + 6 + st];
      d22 = stp[T5 + 12 + st];
      in1 = stp[T5 + 6 + st - 1];
      in2 = stp[T5 + 12 + st - 1];
    }
  }
  __syncwarp();
  const int fp = p.fperiod;
  for (int f = f0; f < f1; ++f) {
    const int64_t gf = foff + f;
    for (int k = lane; k < NM; k += 32) {  // frame prologue (mod.rs:116-125)
      const double cc = d.b[gf * NM + k];
      const double c = f == 0 ? (d.b_first ? d.b_first[(int64_t)u * NM + k] : cc) : s_c[k];  // is_first (mod.rs:79-88)
      s_c[k] = c;
      s_cc[k] = cc;
      s_cinc[k] = (cc - c) / (double)fp;
    }
    __syncwarp();
    double* wv = d.wave + gf * fp;
    for (int i0 = 0; i0 < fp; i0 += 32) {
      const int nch = min(32, fp - i0);
      const double xin = lane < nch ? wv[i0 + lane] : 0.0;
      double xout = 0.0;
      for (int j = 0; j < nch; ++j) {
        double x = __shfl_sync(0xffffffffu, xin, j);
        d11 = iaa * in1 + a * d11;  // df1 (mlsa.rs:54-66)
        d12 = d11 * s_c[1];
        double v = d12 * pp;
        double out = 0.0;
#pragma unroll
        for (int i = 5; i >= 1; --i) {
          const double vi = __shfl_sync(0xffffffffu, v, i - 1);
          x += (i & 1) ? vi : -vi;
          out += vi;
        }
        {
          const double up = __shfl_up_sync(0xffffffffu, d12, 1);
          in1 = (lane == 0) ? x : up;
        }
        x += out;
        d22 = stage_lane ? fir_smem(dl, NM, in2, a, iaa, s_c) : 0.0;  // df2 (mlsa.rs:69-79)
        v = d22 * pp;
        out = 0.0;
#pragma unroll
        for (int i = 5; i >= 1; --i) {
          const double vi = __shfl_sync(0xffffffffu, v, i - 1);
          x += (i & 1) ? vi : -vi;
          out += vi;
        }
        {
          const double up = __shfl_up_sync(0xffffffffu, d22, 1);
          in2 = (lane == 0) ? x : up;
        }
        x += out;
        __syncwarp();
        for (int k = lane; k < NM; k += 32) s_c[k] += s_cinc[k];  // mod.rs:133-135
        __syncwarp();
        if (lane == j) xout = x * p.volume;
      }
      if (lane < nch) wv[i0 + lane] = xout;
    }
    for (int k = lane; k < NM; k += 32) s_c[k] = s_cc[k];  // *coefficients = cc (mod.rs:140)
    __syncwarp();
  }
  if (stp != nullptr) {
    for (int k = lane; k < T5; k += 32) stp[k] = s_dl[k];
    for (int k = lane; k < NM; k += 32) stp[T5 + 18 + k] = s_c[k];
    if (stage_lane) {
      stp[T5 + st] = d11;
      stp[T5 + 6 + st] = d12;
      stp[T5 + 12 + st] = d22;
      if (lane == 0) {
        stp[T5 + 6] = in1;
        stp[T5 + 12] = in2;
      }
    }
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

namespace {
// 16-bit PCM the way the reference's examples write WAV files (examples/genji/main.rs:34-35):
// value.min(i16::MAX).max(i16::MIN) as i16, i.e. clamp, then truncate toward zero. Four samples per
// thread (n and both pointers are multiples of 4 samples: offsets are whole frames).
__global__ void __launch_bounds__(256) k_pcm16(const double* __restrict__ in, int16_t* __restrict__ out, int64_t n4) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  const double2 a = reinterpret_cast<const double2*>(in)[2 * i];
  const double2 c = reinterpret_cast<const double2*>(in)[2 * i + 1];
  auto q = [](double v) -> short { return (short)(int)fmax(fmin(v, 32767.0), -32768.0); };
  short4 o;
  o.x = q(a.x); o.y = q(a.y); o.z = q(c.x); o.w = q(c.y);
  reinterpret_cast<short4*>(out)[i] = o;
}
}  // namespace

int launch_pcm16(const double* in, int16_t* out, int64_t n, cudaStream_t stream) {
  if (n <= 0) return 0;
  if (n % 4 != 0 || ((uintptr_t)in & 15) || ((uintptr_t)out & 7)) return (int)cudaErrorInvalidValue;
  const int64_t n4 = n / 4;
  k_pcm16<<<(unsigned)((n4 + 255) / 256), 256, 0, stream>>>(in, out, n4);
  return (int)cudaGetLastError();
}

int launch_gather_chunk(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d,
                        double* chunk, cudaStream_t stream) {
  if (b.n_utt == 0 || d.frame_count <= 0) return 0;
  const int64_t n = (int64_t)d.frame_count * p.fperiod;
  dim3 grid((unsigned)std::min<int64_t>((n + 255) / 256, 64), (unsigned)b.n_utt);
  k_gather_chunk<<<grid, 256, 0, stream>>>(b, p, d, chunk);
  return (int)cudaGetLastError();
}

int launch_frame_prep(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d,
                      int64_t total_frames, int what, cudaStream_t stream) {
  if (total_frames == 0) return 0;
  const int threads = 128;
  const int blocks = (int)((total_frames + threads - 1) / threads);
  k_frame_prep<<<blocks, threads, 0, stream>>>(p, d, total_frames, what);
  return (int)cudaGetLastError();
}

bool coef_rows_supported(const VocoderParams& p) {
  return p.stage == 0 && p.nmcp >= 2 && (size_t)(kCoefFrames + 1) * (p.nmcp | 1) * 8 <= 48 * 1024;
}
// d.b and d.binc from d.mcp in one kernel: launch_frame_prep(what = 2) + launch_row_increments without b_first.
int launch_coef_rows(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, int64_t total_frames,
                     cudaStream_t stream) {
  if (total_frames == 0) return 0;
  if (d.binc == nullptr || d.b_first != nullptr || !coef_rows_supported(p)) return (int)cudaErrorInvalidValue;
  const size_t smem = (size_t)(kCoefFrames + 1) * (p.nmcp | 1) * 8;
  k_coef_rows<<<(unsigned)((total_frames + kCoefFrames - 1) / kCoefFrames), kCoefFrames, smem, stream>>>(b, p, d,
                                                                                                         total_frames);
  return (int)cudaGetLastError();
}

int launch_pulse(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, int64_t total_samples,
                 cudaStream_t stream) {
  if (b.n_utt == 0) return 0;
  const int threads = 32;
  // whole-batch synthesis: zero the array with one memset and write only the pulses
  const bool whole = d.frame_begin == 0 && d.frame_count < 0;
  if (whole) {
    cudaError_t e = cudaMemsetAsync(d.pulse, 0, (size_t)total_samples * sizeof(double), stream);
    if (e != cudaSuccess) return (int)e;
  }
  if (whole && !d.state) {
    const int64_t total_frames = total_samples / p.fperiod;
    if (total_frames == 0) return 0;
    if (d.runs == nullptr) return (int)cudaErrorInvalidValue;
    cudaError_t e = cudaMemsetAsync(d.runs, 0, sizeof(int32_t), stream);
    if (e != cudaSuccess) return (int)e;
    // at most every other frame starts a run
    const int64_t max_runs = total_frames / 2 + b.n_utt + 1;
    if (p.fperiod % 8 == 0 && p.fperiod <= 8 * kPwMaxBlocks) {
      int* next = d.runs + total_frames + 1;  // the spare slot behind the list (the arena holds total_frames + 2 ints)
      e = cudaMemsetAsync(next, 0, sizeof(int32_t), stream);
      if (e != cudaSuccess) return (int)e;
      k_pulse_list<<<(unsigned)((total_frames + 255) / 256), 256, 0, stream>>>(b, d, total_frames);
      int dev = 0, sms = 148;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      // six blocks per SM: with two the bench batch gains 0.05 ms of the MLPG phase (less crowding of the MCP kernels)
      // but a batch of twice the runs (config 4: 2048 utterances of a higher voice) loses 0.7 ms to the walks
      constexpr int per_sm = 6;
      const int64_t blocks = std::min<int64_t>((max_runs + kPwWarps - 1) / kPwWarps, (int64_t)sms * per_sm);
      k_pulse_warp<<<(unsigned)blocks, 32 * kPwWarps, 0, stream>>>(b, p, d, next);
      return (int)cudaGetLastError();
    }
    k_pulse_list<<<(unsigned)((total_frames + 255) / 256), 256, 0, stream>>>(b, d, total_frames);
    const int64_t max_threads = max_runs * (32 / kRunLanes);
    k_pulse_runs<<<(unsigned)((max_threads + 127) / 128), 128, 0, stream>>>(b, p, d, total_frames);
    return (int)cudaGetLastError();
  }
  k_pulse<<<(b.n_utt + threads - 1) / threads, threads, 0, stream>>>(b, p, d, whole ? 0 : 1);
  return (int)cudaGetLastError();
}

int launch_excitation(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d,
                      int64_t total_frames, cudaStream_t stream) {
  if (total_frames == 0) return 0;
  if (p.nlpf > kMaxLpf || p.fperiod > kMaxFperiod) return (int)cudaErrorInvalidValue;
  if (p.nlpf == 31 && p.fperiod % 8 == 0 && p.fperiod <= 256)
  {
    const int64_t items = d.frame_count >= 0 ? (int64_t)b.n_utt * d.frame_count : total_frames;
    if (items == 0) return 0;
    k_excitation_w<31><<<(unsigned)((items + kExWarps * kExGroup - 1) / (kExWarps * kExGroup)), 32 * kExWarps, 0,
                         stream>>>(b, p, d, total_frames);
  }
  else
    k_excitation<<<(unsigned)total_frames, 256, 0, stream>>>(b, p, d, total_frames);
  return (int)cudaGetLastError();
}

int launch_mlsa(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, int variant,
                cudaStream_t stream) {
  if (b.n_utt == 0) return 0;
  if (p.nmcp != 35 || (variant != 0 && (!(p.alpha > 1e-3) || p.fperiod % 16 != 0))) {
    // no register-resident kernel for this vector length / frame period (or alpha ~ 0: the blocked scan divides by
    // it): the generic one. Streaming state fits the per-utterance slot up to nmcp = 35.
    if (p.nmcp < 2 || p.nmcp > kAnyMaxNM || (d.state != nullptr && p.nmcp > 35)) return (int)cudaErrorInvalidValue;
    const size_t smem = (size_t)(5 * (p.nmcp - 1) + 3 * p.nmcp) * sizeof(double);
    k_mlsa_any<<<b.n_utt, 32, smem, stream>>>(b, p, d);
    return (int)cudaGetLastError();
  }
  // variants 1 / 2 (default): blocked-scan kernel, one warp per utterance (mlsa.cu)
  if (variant != 0) return launch_mlsa_v2(b, p, d, stream);
  constexpr int G = 8, WPB = 1;
  const int upb = (32 / G) * WPB;
  k_mlsa_v0<G, 35, WPB><<<(b.n_utt + upb - 1) / upb, 32 * WPB, 0, stream>>>(b, p, d);
  return (int)cudaGetLastError();
}

}  // namespace jbk
