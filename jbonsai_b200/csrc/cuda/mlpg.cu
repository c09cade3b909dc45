// MLPG kernel (sm_100a): banded W'U^-1W build, LDL^T factorisation, forward/backward
// substitution and the GV Newton adjustment — one independent system per
// (utterance, stream, dimension), one thread per system, time recurrences in-thread.
//
// Reference path: MlpgAdjust::create (src/mlpg_adjust/mod.rs:51-95), Mask (mask.rs:18-83),
// MlpgMatrix / MlpgGlobalVariance (mlpg.rs:22-292), MeanVari::with_ivar (model/mean_vari.rs:21-31).
//
// Layout: every per-frame work plane is [frame][dim] so that the threads of a warp
// (consecutive dims of one utterance) touch consecutive doubles. Compile with -fmad=false:
// the operation order below is the reference's and no multiply-add may be contracted, which
// makes the generated parameters bit-identical to the CPU restatement.
#include <climits>

#include <cstdlib>

#include "kernels.h"

namespace jbk {

namespace {

constexpr double kNoData = -1e10;

// ---------------------------------------------------------------- expand
// One CTA per utterance: duration expansion (IterExt::duration, mod.rs:105-112), MSD mask
// (Mask::create, mask.rs:20-28), boundary distances (mask.rs:51-82) and voiced compaction, as
// block-wide scans over tiles of kExpandThreads states / frames (sum, running max, running min).
constexpr int kExpandThreads = 256;

// Inclusive block scan of (sum of a, max of b) in tile order; returns the scanned pair of this
// thread and leaves the tile totals in tot[0], tot[1]. sw: 2 * (kExpandThreads / 32) ints.
__device__ __forceinline__ void block_scan_sum_max(int& a, int& b, int* sw, int* tot) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int ya = __shfl_up_sync(0xffffffffu, a, o), yb = __shfl_up_sync(0xffffffffu, b, o);
    if (lane >= o) { a += ya; b = max(b, yb); }
  }
  if (lane == 31) { sw[2 * w] = a; sw[2 * w + 1] = b; }
  __syncthreads();
  int pa = 0, pb = INT_MIN;
  for (int i = 0; i < w; ++i) { pa += sw[2 * i]; pb = max(pb, sw[2 * i + 1]); }
  a += pa;
  b = max(b, pb);
  if (threadIdx.x == kExpandThreads - 1) { tot[0] = a; tot[1] = b; }
  __syncthreads();
}

__global__ void __launch_bounds__(kExpandThreads) k_expand(BatchLayout b, StreamSet set) {
  __shared__ int sw[2 * (kExpandThreads / 32)];
  __shared__ int tot[2];
  const int u = blockIdx.x, tid = threadIdx.x;
  const int64_t soff = b.state_off[u], foff = b.frame_off[u];
  const int S = (int)(b.state_off[u + 1] - soff);
  const int T = (int)(b.frame_off[u + 1] - foff);
  // frame -> state: exclusive prefix sum of the durations, every state writes its own run
  int carry = 0;
  for (int s0 = 0; s0 < S; s0 += kExpandThreads) {
    const int s = s0 + tid;
    const int dur = s < S ? b.durations[soff + s] : 0;
    int a = dur, dummy = 0;
    block_scan_sum_max(a, dummy, sw, tot);
    const int start = carry + a - dur;
    for (int k = 0; k < dur; ++k) b.frame_state[foff + start + k] = s;
    carry += tot[0];
    __syncthreads();
  }
  __syncthreads();
  for (int si = 0; si < set.n; ++si) {
    const StreamDev& sd = set.s[si];
    if (!sd.is_msd) continue;
    // forward: voiced count (compaction index) and last masked frame so far
    int cnt = 0, last = -1;
    for (int f0 = 0; f0 < T; f0 += kExpandThreads) {
      const int f = f0 + tid;
      const bool m = f < T && sd.msd[soff + b.frame_state[foff + f]] > sd.msd_threshold;
      int a = m ? 1 : 0, bb = (f < T && !m) ? f : -1;
      block_scan_sum_max(a, bb, sw, tot);
      if (f < T) {
        if (m) {
          sd.left[foff + f] = f - (max(last, bb) + 1);
          sd.vidx[foff + cnt + a - 1] = f;
        } else {
          sd.left[foff + f] = -1;  // masked
        }
      }
      cnt += tot[0];
      last = max(last, tot[1]);
      __syncthreads();
    }
    if (tid == 0) sd.nvoiced[u] = cnt;
    // backward: first masked frame at or after f, as a running max of the negated index
    int next = T;
    for (int f0 = 0; f0 < T; f0 += kExpandThreads) {
      const int f = T - 1 - (f0 + tid);
      const bool m = f >= 0 && sd.left[foff + f] >= 0;
      int a = 0, bb = (f >= 0 && !m) ? -f : INT_MIN;
      block_scan_sum_max(a, bb, sw, tot);
      if (f >= 0) sd.right[foff + f] = m ? (min(next, bb == INT_MIN ? INT_MAX : -bb) - 1 - f) : -1;
      if (tot[1] != INT_MIN) next = min(next, -tot[1]);
      __syncthreads();
    }
  }
}

__device__ __forceinline__ double with_ivar(double vari) {
  double a = fabs(vari);
  if (a > 1e19) return 0.0;
  if (a < 1e-19) return 1e38;
  return 1.0 / vari;
}

// ---------------------------------------------------------------- solve (+GV)
template <int BW>
__global__ void __launch_bounds__(128) k_mlpg(BatchLayout b, StreamDev s, int64_t plane) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int u = gid / s.dim, d = gid - u * s.dim;
  if (u >= b.n_utt) return;
  const int64_t soff = b.state_off[u], foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  const int dim = s.dim, L = s.dim * s.nwin;
  const int Tp = s.is_msd ? s.nvoiced[u] : T;
  const bool msd = s.is_msd != 0;

  // scratch planes, all indexed by compacted frame k
  double* const base = s.scratch + foff * dim + d;
  auto F = [&](int j, int k) -> double& { return base[(int64_t)j * plane + (int64_t)k * dim]; };  // factorised band
  double* const pwum = base + (int64_t)BW * plane;
  double* const pg = base + (int64_t)(BW + 1) * plane;  // g, then par
  auto O = [&](int j, int k) -> double& {  // un-factorised band (GV only)
    return base[(int64_t)(BW + 2 + j) * plane + (int64_t)k * dim];
  };
  double* const pgrad = base + (int64_t)(2 * BW + 2) * plane;

  auto frame_of = [&](int k) -> int { return msd ? s.vidx[foff + k] : k; };

  // (mean, ivar) of window w at compacted index k  (mod.rs:57-84)
  auto param = [&](int w, int k, double& mean, double& ivar) {
    const int t = frame_of(k);
    const int st = b.frame_state[foff + t];
    const int64_t o = (soff + st) * L + dim * w + d;
    mean = s.mean[o];
    ivar = with_ivar(s.vari[o]);
    if (w != 0) {
      const int l = msd ? s.left[foff + t] : t;
      const int r = msd ? s.right[foff + t] : T - 1 - t;
      const int wd = s.windows.width[w];
      const int lw = wd / 2, rw = wd - lw - 1;
      if (l < lw || r < rw) ivar = 0.0;
    }
  };

  // ---- pass 1: band build (mlpg.rs:31-61), LDL (mlpg.rs:79-92), forward substitution (:98-104)
  double prev[BW][BW];  // prev[i] = factorised row k-i (i >= 1)
  double gprev[BW];
#pragma unroll
  for (int i = 0; i < BW; ++i) {
    gprev[i] = 0.0;
#pragma unroll
    for (int j = 0; j < BW; ++j) prev[i][j] = 0.0;
  }
  for (int k = 0; k < Tp; ++k) {
    double row[BW];
#pragma unroll
    for (int j = 0; j < BW; ++j) row[j] = 0.0;
    double wum = 0.0;
    for (int w = 0; w < s.nwin; ++w) {
      const int wd = s.windows.width[w];
      for (int index = wd - 1; index >= 0; --index) {
        const double coef = s.windows.coef[w][index];
        if (coef == 0.0) continue;
        const int idx = k - (index - wd / 2);
        if (idx < 0 || idx >= Tp) continue;
        double mean, ivar;
        param(w, idx, mean, ivar);
        const double wu = coef * ivar;
        wum += wu * mean;
        for (int inner = wd - 1; inner >= index; --inner) {
          const double c2 = s.windows.coef[w][inner];
          if (c2 == 0.0) continue;
          const int j = inner - index;
          if (k + j >= Tp) break;
#pragma unroll
          for (int jj = 0; jj < BW; ++jj)
            if (jj == j) row[jj] += wu * c2;
        }
      }
    }
    if (s.has_gv) {
#pragma unroll
      for (int j = 0; j < BW; ++j) O(j, k) = row[j];
    }
    pwum[(int64_t)k * dim] = wum;
#pragma unroll
    for (int i = 1; i < BW; ++i)
      if (i < k + 1) row[0] -= prev[i][i] * prev[i][i] * prev[i][0];
#pragma unroll
    for (int i = 1; i < BW; ++i) {
#pragma unroll
      for (int j = 1; j < BW; ++j)
        if (j < BW - i && j < k + 1) row[i] -= prev[j][j] * prev[j][i + j] * prev[j][0];
      row[i] /= row[0];
    }
    double g = wum;
#pragma unroll
    for (int i = 1; i < BW; ++i)
      if (i < k + 1) g -= prev[i][i] * gprev[i];
#pragma unroll
    for (int j = 0; j < BW; ++j) F(j, k) = row[j];
    pg[(int64_t)k * dim] = g;
#pragma unroll
    for (int i = BW - 1; i >= 2; --i) {
      gprev[i] = gprev[i - 1];
#pragma unroll
      for (int j = 0; j < BW; ++j) prev[i][j] = prev[i - 1][j];
    }
    if (BW > 1) {
      gprev[1] = g;
#pragma unroll
      for (int j = 0; j < BW; ++j) prev[1][j] = row[j];
    }
  }

  // ---- pass 2: backward substitution (mlpg.rs:106-113); par overwrites g in place
  {
    double pn[BW];  // pn[i] = par[k+i]
#pragma unroll
    for (int i = 0; i < BW; ++i) pn[i] = 0.0;
    for (int k = Tp - 1; k >= 0; --k) {
      double par = pg[(int64_t)k * dim] / F(0, k);
#pragma unroll
      for (int i = 1; i < BW; ++i)
        if (i < Tp - k) par -= F(i, k) * pn[i];
      pg[(int64_t)k * dim] = par;
#pragma unroll
      for (int i = BW - 1; i >= 2; --i) pn[i] = pn[i - 1];
      if (BW > 1) pn[1] = par;
    }
  }
  double* const ppar = pg;
  auto PAR = [&](int k) -> double& { return ppar[(int64_t)k * dim]; };

  // ---- GV (MlpgGlobalVariance::parmgen, mlpg.rs:258-292), reductions in increasing k
  if (s.has_gv && Tp > 0) {
    auto sw = [&](int k) -> bool { return s.gv_switch[soff + b.frame_state[foff + frame_of(k)]] != 0; };
    int gv_length = 0;
    for (int k = 0; k < Tp; ++k) gv_length += sw(k) ? 1 : 0;
    if (gv_length > 0) {
      const double gv_mean = s.gv_mean[(int64_t)u * dim + d] * s.gv_weight;  // mlpg.rs:137-138
      const double gv_vari = s.gv_vari[(int64_t)u * dim + d];
      auto calc_gv = [&](double& mean, double& vari) {  // mlpg.rs:168-188
        double sum = 0.0;
        for (int k = 0; k < Tp; ++k)
          if (sw(k)) sum += PAR(k);
        mean = sum / (double)gv_length;
        double acc = 0.0;
        for (int k = 0; k < Tp; ++k)
          if (sw(k)) {
            const double p = PAR(k);
            acc += (p - mean) * (p - mean);
          }
        vari = acc / (double)gv_length;
      };
      {  // conv_gv (mlpg.rs:191-199)
        double mean, vari;
        calc_gv(mean, vari);
        const double ratio = sqrt(gv_mean / vari);
        for (int k = 0; k < Tp; ++k)
          if (sw(k)) PAR(k) = ratio * (PAR(k) - mean) + mean;
      }
      const double wnorm = 1.0 / (double)((int64_t)s.nwin * (int64_t)Tp);
      const double len = (double)Tp;
      const double len2 = (double)((int64_t)Tp * (int64_t)Tp);
      const double lenm1 = (double)(Tp - 1);
      double step = 0.1, prevobj = 0.0;
      for (int it = 1; it <= 5; ++it) {
        double mean, vari;
        calc_gv(mean, vari);
        const double gvobj = -0.5 * 1.0 * vari * gv_vari * (vari - 2.0 * gv_mean);
        // calc_hmmobj_derivative (mlpg.rs:200-226)
        double hmmobj = 0.0;
        for (int k = 0; k < Tp; ++k) {
          double g = O(0, k) * PAR(k);
#pragma unroll
          for (int i = 1; i < BW; ++i) {
            if (k + i < Tp) g += O(i, k) * PAR(k + i);
            if (k + 1 > i) g += O(i, k - i) * PAR(k - i);
          }
          pgrad[(int64_t)k * dim] = g;
        }
        for (int k = 0; k < Tp; ++k)
          hmmobj += 1.0 * wnorm * PAR(k) * (pwum[(int64_t)k * dim] - 0.5 * pgrad[(int64_t)k * dim]);
        const double obj = -(hmmobj + gvobj);
        if (it > 1) {
          if (obj > prevobj)
            step *= 0.5;
          else if (obj < prevobj)
            step *= 1.2;
        }
        // next_step (mlpg.rs:227-256)
        const double dv = -2.0 * gv_vari * (vari - gv_mean) / len;
        for (int k = 0; k < Tp; ++k) {
          const double p = PAR(k);
          const double h = -1.0 * wnorm * O(0, k) -
                           1.0 * 2.0 / len2 *
                               (lenm1 * gv_vari * (vari - gv_mean) +
                                2.0 * gv_vari * (p - mean) * (p - mean));
          const double gk = pgrad[(int64_t)k * dim], wk = pwum[(int64_t)k * dim];
          double next_g;
          if (sw(k))
            next_g = 1.0 / h * (1.0 * wnorm * (-gk + wk) + 1.0 * dv * (p - mean));
          else
            next_g = 1.0 / h * (1.0 * wnorm * (-gk + wk));
          PAR(k) = p + step * next_g;
        }
        prevobj = obj;
      }
    }
  }

  // ---- Mask::fill (mask.rs:34-49): scatter back, NODATA elsewhere
  double* out = s.par + foff * dim + d;
  if (msd) {
    for (int t = 0; t < T; ++t)
      if (s.left[foff + t] < 0) out[(int64_t)t * dim] = kNoData;
    for (int k = 0; k < Tp; ++k) out[(int64_t)s.vidx[foff + k] * dim] = PAR(k);
  } else {
    for (int k = 0; k < Tp; ++k) out[(int64_t)k * dim] = PAR(k);
  }
}


// ================================================================ fast path
// Same arithmetic as k_mlpg above (reference operation order for the band build, the LDL^T
// factorisation and both substitutions), restructured for the memory system:
//   * the (mean, ivar) pair of a frame only changes when the frame's state changes, so each
//     thread keeps the current state's values in registers (one reciprocal per state and window,
//     not seven per frame) and slides a 3-frame window over them;
//   * the GV adjustment runs in its own kernel (k_mlpg_gv), one CTA per system with the system
//     in shared memory, data-parallel over frames.
// Window shapes: kShape1 = one static window of width 1 (band width 1); kShape133 = static + two
// 3-tap dynamic windows (band width 3) — every shipped voice. Other shapes use k_mlpg.
constexpr int kShape1 = 0, kShape133 = 1;

template <int SHAPE>
__global__ void __launch_bounds__(128) k_mlpg_solve(BatchLayout b, StreamDev s, int64_t plane) {
  constexpr int BW = SHAPE == kShape133 ? 3 : 1;
  constexpr int NW = SHAPE == kShape133 ? 3 : 1;
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int u = gid / s.dim, d = gid - u * s.dim;
  if (u >= b.n_utt) return;
  const int64_t soff = b.state_off[u], foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  const int dim = s.dim, L = s.dim * s.nwin;
  const bool msd = s.is_msd != 0;
  const int Tp = msd ? s.nvoiced[u] : T;

  double* const base = s.scratch + foff * dim + d;
  double* const pwum = base + (int64_t)BW * plane;
  double* const pg = base + (int64_t)(BW + 1) * plane;
  const bool gv = s.has_gv != 0;
  const int32_t* const fstate = b.frame_state + foff;
  const int32_t* const vidx = msd ? s.vidx + foff : nullptr;
  const int32_t* const left = msd ? s.left + foff : nullptr;
  const int32_t* const right = msd ? s.right + foff : nullptr;

  // window coefficients (uniform)
  double wc[NW][3];
#pragma unroll
  for (int w = 0; w < NW; ++w)
#pragma unroll
    for (int i = 0; i < 3; ++i) wc[w][i] = (w == 0) ? (i == 0 ? s.windows.coef[0][0] : 0.0) : s.windows.coef[w][i];

  // (mean, ivar) of the current state, per window; refreshed when the state index changes
  int st_cached = -1;
  double cm[NW], civ[NW];
  // sliding window over compacted frames k-1, k, k+1: effective (mean, ivar) per window
  double m_[NW][3], iv_[NW][3];
#pragma unroll
  for (int w = 0; w < NW; ++w)
#pragma unroll
    for (int i = 0; i < 3; ++i) { m_[w][i] = 0.0; iv_[w][i] = 0.0; }

  auto fetch = [&](int k, double (&mo)[NW], double (&ivo)[NW]) {  // params of compacted frame k (mod.rs:57-84)
    const int t = msd ? vidx[k] : k;
    const int st = fstate[t];
    if (st != st_cached) {
      st_cached = st;
      const int64_t o = (soff + st) * L + d;
#pragma unroll
      for (int w = 0; w < NW; ++w) {
        cm[w] = s.mean[o + dim * w];
        civ[w] = with_ivar(s.vari[o + dim * w]);
      }
    }
    const int l = msd ? left[t] : t;
    const int r = msd ? right[t] : T - 1 - t;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      mo[w] = cm[w];
      ivo[w] = (w != 0 && (l < 1 || r < 1)) ? 0.0 : civ[w];  // 3-tap window: left/right width 1
    }
  };

  // ---- pass 1: band build (mlpg.rs:31-61), LDL (mlpg.rs:79-92), forward substitution (:98-104)
  double prev[BW][BW], gprev[BW];
#pragma unroll
  for (int i = 0; i < BW; ++i) {
    gprev[i] = 0.0;
#pragma unroll
    for (int j = 0; j < BW; ++j) prev[i][j] = 0.0;
  }
  if (Tp > 0) {
    double mo[NW], ivo[NW];
    fetch(0, mo, ivo);
#pragma unroll
    for (int w = 0; w < NW; ++w) { m_[w][1] = mo[w]; iv_[w][1] = ivo[w]; }  // slot 1 = frame k, slot 2 = k+1
    if (Tp > 1) {
      fetch(1, mo, ivo);
#pragma unroll
      for (int w = 0; w < NW; ++w) { m_[w][2] = mo[w]; iv_[w][2] = ivo[w]; }
    }
  }
  for (int k = 0; k < Tp; ++k) {
    double row[BW];
#pragma unroll
    for (int j = 0; j < BW; ++j) row[j] = 0.0;
    double wum = 0.0;
    {  // window 0: static, width 1
      const double coef = wc[0][0];
      if (coef != 0.0) {
        const double wu = coef * iv_[0][1];
        wum += wu * m_[0][1];
        row[0] += wu * coef;
      }
    }
    if (SHAPE == kShape133) {
#pragma unroll
      for (int w = 1; w < NW; ++w) {
#pragma unroll
        for (int index = 2; index >= 0; --index) {
          const double coef = wc[w][index];
          if (coef == 0.0) continue;
          const int idx = k - (index - 1);  // slot = 1 - (index - 1) = 2 - index
          if (idx < 0 || idx >= Tp) continue;
          const double wu = coef * iv_[w][2 - index];
          wum += wu * m_[w][2 - index];
#pragma unroll
          for (int inner = 2; inner >= index; --inner) {
            const double c2 = wc[w][inner];
            if (c2 == 0.0) continue;
            const int j = inner - index;
            if (k + j >= Tp) break;
            row[j] += wu * c2;
          }
        }
      }
    }
    if (gv) {
#pragma unroll
      for (int j = 0; j < BW; ++j) base[(int64_t)(BW + 2 + j) * plane + (int64_t)k * dim] = row[j];
    }
    pwum[(int64_t)k * dim] = wum;
#pragma unroll
    for (int i = 1; i < BW; ++i)
      if (i < k + 1) row[0] -= prev[i][i] * prev[i][i] * prev[i][0];
#pragma unroll
    for (int i = 1; i < BW; ++i) {
#pragma unroll
      for (int j = 1; j < BW; ++j)
        if (j < BW - i && j < k + 1) row[i] -= prev[j][j] * prev[j][i + j] * prev[j][0];
      row[i] /= row[0];
    }
    double g = wum;
#pragma unroll
    for (int i = 1; i < BW; ++i)
      if (i < k + 1) g -= prev[i][i] * gprev[i];
#pragma unroll
    for (int j = 0; j < BW; ++j) base[(int64_t)j * plane + (int64_t)k * dim] = row[j];
    pg[(int64_t)k * dim] = g;
#pragma unroll
    for (int i = BW - 1; i >= 2; --i) {
      gprev[i] = gprev[i - 1];
#pragma unroll
      for (int j = 0; j < BW; ++j) prev[i][j] = prev[i - 1][j];
    }
    if (BW > 1) {
      gprev[1] = g;
#pragma unroll
      for (int j = 0; j < BW; ++j) prev[1][j] = row[j];
    }
    // slide the parameter window
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      m_[w][0] = m_[w][1]; iv_[w][0] = iv_[w][1];
      m_[w][1] = m_[w][2]; iv_[w][1] = iv_[w][2];
    }
    if (k + 2 < Tp) {
      double mo[NW], ivo[NW];
      fetch(k + 2, mo, ivo);
#pragma unroll
      for (int w = 0; w < NW; ++w) { m_[w][2] = mo[w]; iv_[w][2] = ivo[w]; }
    }
  }

  // ---- pass 2: backward substitution (mlpg.rs:106-113); loads of the next steps are issued ahead
  double* out = s.par + foff * dim + d;
  const bool direct = !gv;  // without GV this kernel also writes the output
  if (direct && msd)
    for (int t = 0; t < T; ++t)
      if (left[t] < 0) out[(int64_t)t * dim] = kNoData;
  {
    double pn[BW];
#pragma unroll
    for (int i = 0; i < BW; ++i) pn[i] = 0.0;
    constexpr int U = 4;
    for (int k0 = Tp - 1; k0 >= 0; k0 -= U) {
      double fg[U], f[U][BW];
#pragma unroll
      for (int q = 0; q < U; ++q) {
        const int k = k0 - q;
        if (k >= 0) {
          fg[q] = pg[(int64_t)k * dim];
#pragma unroll
          for (int j = 0; j < BW; ++j) f[q][j] = base[(int64_t)j * plane + (int64_t)k * dim];
        }
      }
#pragma unroll
      for (int q = 0; q < U; ++q) {
        const int k = k0 - q;
        if (k < 0) break;
        double par = fg[q] / f[q][0];
#pragma unroll
        for (int i = 1; i < BW; ++i)
          if (i < Tp - k) par -= f[q][i] * pn[i];
        pg[(int64_t)k * dim] = par;
        if (direct) out[(int64_t)(msd ? vidx[k] : k) * dim] = par;
#pragma unroll
        for (int i = BW - 1; i >= 2; --i) pn[i] = pn[i - 1];
        if (BW > 1) pn[1] = par;
      }
    }
  }
}


// ================================================================ split path (shape 1-3-3, the shipped voices)
// The band build of mlpg.rs:31-61 has no dependence along time: frame k only needs the (mean, ivar)
// pairs of frames k-1, k, k+1. Only the LDL^T factorisation and the two substitutions are serial. So:
//   k_state_ivar   ivar = MeanVari::with_ivar(vari), one division per (state, window, dim)
//   k_band_build   thread per compacted frame (loops over dim): un-factorised band + W'U^-1 mu,
//                  written "system-major" ([utt][dim][frame]: the T frames of one system contiguous)
//   k_ldl          thread per system: factorisation + forward substitution reading its own
//                  contiguous stream four frames ahead (no data-dependent addresses, so the loads
//                  run ahead of the recurrence), factorised rows written [frame][dim] (coalesced
//                  over the dims of an utterance), then the backward substitution
//   k_mlpg_gv      CTA per system, reads the system-major planes with unit stride
// Every floating-point operation and its order is the one of k_mlpg_solve / the reference; only
// where it runs changed, so the parameters stay bit-identical.
// Plane map (each plane = total_frames * dim doubles): [1, BW) off-diagonals of the factorised band (plane 0, its
// diagonal, is not stored: k_ldl keeps g / diagonal instead of g), BW: g / diagonal, then
// system-major: BW+1: W'U^-1 mu, BW+2+j: un-factorised band j, 2BW+2: solution (GV input),
// 2BW+3: GV gradient (only when a system does not fit in shared memory).
__global__ void __launch_bounds__(256) k_state_ivar(const double* __restrict__ vari, double* __restrict__ ivar,
                                                    int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) ivar[i] = with_ivar(vari[i]);
}

__global__ void __launch_bounds__(128) k_band_build(BatchLayout b, StreamDev s, int64_t plane) {
  constexpr int BW = 3, NW = 3;
  const int u = blockIdx.y;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t soff = b.state_off[u], foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  const bool msd = s.is_msd != 0;
  const int Tp = msd ? s.nvoiced[u] : T;
  if (k >= Tp) return;
  const int dim = s.dim, L = s.dim * s.nwin;
  const int32_t* const fstate = b.frame_state + foff;

  double wc[NW][3];
#pragma unroll
  for (int w = 0; w < NW; ++w)
#pragma unroll
    for (int i = 0; i < 3; ++i) wc[w][i] = (w == 0) ? (i == 0 ? s.windows.coef[0][0] : 0.0) : s.windows.coef[w][i];

  // compacted frames k-1, k, k+1: state row and "inside the mask" flag (mod.rs:57-84)
  int64_t srow[3];
  bool edge[3], valid[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int kk = k - 1 + i;
    valid[i] = kk >= 0 && kk < Tp;
    const int t = valid[i] ? (msd ? s.vidx[foff + kk] : kk) : 0;
    srow[i] = (soff + fstate[t]) * L;
    const int l = msd ? s.left[foff + t] : t;
    const int r = msd ? s.right[foff + t] : T - 1 - t;
    edge[i] = l < 1 || r < 1;  // 3-tap window: left/right width 1
  }
  double* const sys = s.scratch + foff * dim + k;  // + d * T: system d of this utterance
  for (int d = 0; d < dim; ++d) {
    double m_[NW][3], iv_[NW][3];
#pragma unroll
    for (int w = 0; w < NW; ++w)
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const bool need = valid[i] && (w != 0 || i == 1);
        m_[w][i] = need ? s.mean[srow[i] + dim * w + d] : 0.0;
        iv_[w][i] = need ? ((w != 0 && edge[i]) ? 0.0 : s.ivar[srow[i] + dim * w + d]) : 0.0;
      }
    double row[BW] = {0.0, 0.0, 0.0};
    double wum = 0.0;
    {  // window 0: static, width 1
      const double coef = wc[0][0];
      if (coef != 0.0) {
        const double wu = coef * iv_[0][1];
        wum += wu * m_[0][1];
        row[0] += wu * coef;
      }
    }
#pragma unroll
    for (int w = 1; w < NW; ++w) {
#pragma unroll
      for (int index = 2; index >= 0; --index) {
        const double coef = wc[w][index];
        if (coef == 0.0) continue;
        const int idx = k - (index - 1);  // slot = 2 - index
        if (idx < 0 || idx >= Tp) continue;
        const double wu = coef * iv_[w][2 - index];
        wum += wu * m_[w][2 - index];
#pragma unroll
        for (int inner = 2; inner >= index; --inner) {
          const double c2 = wc[w][inner];
          if (c2 == 0.0) continue;
          const int j = inner - index;
          if (k + j >= Tp) break;
          row[j] += wu * c2;
        }
      }
    }
    double* o = sys + (int64_t)d * T;
    o[(int64_t)(BW + 1) * plane] = wum;
#pragma unroll
    for (int j = 0; j < BW; ++j) o[(int64_t)(BW + 2 + j) * plane] = row[j];
  }
}

__global__ void __launch_bounds__(128) k_ldl(BatchLayout b, StreamDev s, int64_t plane) {
  constexpr int BW = 3, U = 8;
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int u = gid / s.dim, d = gid - u * s.dim;
  if (u >= b.n_utt) return;
  const int64_t foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  const int dim = s.dim;
  const bool msd = s.is_msd != 0;
  const int Tp = msd ? s.nvoiced[u] : T;
  const bool gv = s.has_gv != 0;

  double* const base = s.scratch + foff * dim + d;             // [frame][dim] planes
  double* const pg = base + (int64_t)BW * plane;
  double* const sys = s.scratch + foff * dim + (int64_t)d * T;  // system-major planes
  const double* const swum = sys + (int64_t)(BW + 1) * plane;
  const double* const sband = sys + (int64_t)(BW + 2) * plane;
  double* const spar = sys + (int64_t)(2 * BW + 2) * plane;

  // ---- LDL (mlpg.rs:79-92) + forward substitution (:98-104)
  double prev[BW][BW], gprev[BW];
#pragma unroll
  for (int i = 0; i < BW; ++i) {
    gprev[i] = 0.0;
#pragma unroll
    for (int j = 0; j < BW; ++j) prev[i][j] = 0.0;
  }
  // the loads of batch k0 + U are in flight while batch k0 is factorised
  double in[U][BW], inw[U], nin[U][BW], ninw[U];
  auto load_fwd = [&](int k0, double (&a)[U][BW], double (&w)[U]) {
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const int k = k0 + q;
      w[q] = 0.0;
#pragma unroll
      for (int j = 0; j < BW; ++j) a[q][j] = 0.0;
      if (k < Tp) {
        w[q] = __ldg(swum + k);
#pragma unroll
        for (int j = 0; j < BW; ++j) a[q][j] = __ldg(sband + (int64_t)j * plane + k);
      }
    }
  };
  load_fwd(0, nin, ninw);
  for (int k0 = 0; k0 < Tp; k0 += U) {
#pragma unroll
    for (int q = 0; q < U; ++q) {
      inw[q] = ninw[q];
#pragma unroll
      for (int j = 0; j < BW; ++j) in[q][j] = nin[q][j];
    }
    load_fwd(k0 + U, nin, ninw);
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const int k = k0 + q;
      if (k >= Tp) break;
      double row[BW];
#pragma unroll
      for (int j = 0; j < BW; ++j) row[j] = in[q][j];
#pragma unroll
      for (int i = 1; i < BW; ++i)
        if (i < k + 1) row[0] -= prev[i][i] * prev[i][i] * prev[i][0];
#pragma unroll
      for (int i = 1; i < BW; ++i) {
#pragma unroll
        for (int j = 1; j < BW; ++j)
          if (j < BW - i && j < k + 1) row[i] -= prev[j][j] * prev[j][i + j] * prev[j][0];
        row[i] /= row[0];
      }
      double g = inw[q];
#pragma unroll
      for (int i = 1; i < BW; ++i)
        if (i < k + 1) g -= prev[i][i] * gprev[i];
      // what the backward pass needs of frame k: g / row[0] (the division of mlpg.rs:108, same operands, done here
      // off the recurrence's chain) and the two off-diagonals: three planes instead of four, both ways
#pragma unroll
      for (int j = 1; j < BW; ++j) base[(int64_t)j * plane + (int64_t)k * dim] = row[j];
      pg[(int64_t)k * dim] = g / row[0];
#pragma unroll
      for (int i = BW - 1; i >= 2; --i) {
        gprev[i] = gprev[i - 1];
#pragma unroll
        for (int j = 0; j < BW; ++j) prev[i][j] = prev[i - 1][j];
      }
      gprev[1] = g;
#pragma unroll
      for (int j = 0; j < BW; ++j) prev[1][j] = row[j];
    }
  }

  // ---- backward substitution (mlpg.rs:106-113)
  double* out = s.par + foff * dim + d;
  const int32_t* const vidx = msd ? s.vidx + foff : nullptr;
  if (!gv && msd)
    for (int t = 0; t < T; ++t)
      if (s.left[foff + t] < 0) out[(int64_t)t * dim] = kNoData;
  double pn[BW];
#pragma unroll
  for (int i = 0; i < BW; ++i) pn[i] = 0.0;
  double fg[U], f[U][BW], nfg[U], nf[U][BW];
  auto load_bwd = [&](int k0, double (&a)[U][BW], double (&w)[U]) {
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const int k = k0 - q;
      w[q] = 0.0;
#pragma unroll
      for (int j = 0; j < BW; ++j) a[q][j] = 0.0;
      if (k >= 0) {
        w[q] = pg[(int64_t)k * dim];
#pragma unroll
        for (int j = 1; j < BW; ++j) a[q][j] = base[(int64_t)j * plane + (int64_t)k * dim];
      }
    }
  };
  load_bwd(Tp - 1, nf, nfg);
  for (int k0 = Tp - 1; k0 >= 0; k0 -= U) {
#pragma unroll
    for (int q = 0; q < U; ++q) {
      fg[q] = nfg[q];
#pragma unroll
      for (int j = 0; j < BW; ++j) f[q][j] = nf[q][j];
    }
    load_bwd(k0 - U, nf, nfg);
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const int k = k0 - q;
      if (k < 0) break;
      double par = fg[q];  // g / row[0], divided in the forward pass
#pragma unroll
      for (int i = 1; i < BW; ++i)
        if (i < Tp - k) par -= f[q][i] * pn[i];
      if (gv) spar[k] = par;
      else out[(int64_t)(msd ? vidx[k] : k) * dim] = par;
#pragma unroll
      for (int i = BW - 1; i >= 2; --i) pn[i] = pn[i - 1];
      pn[1] = par;
    }
  }
}

// Static window only (band width 1), no mask, no GV — the LPF stream: every frame is its own
// 1 x 1 system, par = (c iv m) / (c iv c) with the operations of mlpg.rs:31-61,106-113.
// A block owns kStaticFrames consecutive frames: one utterance search and one state lookup per FRAME (a thread per
// element repeated both for every dimension: 0.36 ms on the bench batch, all of it on top of the MCP chain it runs
// beside), then the block walks the frames' elements in order, so the output is written coalesced.
constexpr int kStaticFrames = 64;
__global__ void __launch_bounds__(256) k_mlpg_static(BatchLayout b, StreamDev s, int64_t total_frames) {
  __shared__ int64_t row_of[kStaticFrames];  // state row (in units of vectors) of every frame of the block
  const int dim = s.dim;
  const int64_t f0 = (int64_t)blockIdx.x * kStaticFrames;
  const int nf = (int)min((int64_t)kStaticFrames, total_frames - f0);
  if ((int)threadIdx.x < nf) {
    const int64_t gf = f0 + threadIdx.x;
    int lo = 0, hi = b.n_utt;  // utterance of this frame
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (b.frame_off[mid] <= gf) lo = mid; else hi = mid;
    }
    row_of[threadIdx.x] = (b.state_off[lo] + b.frame_state[gf]) * (int64_t)(dim * s.nwin);
  }
  __syncthreads();
  const double coef = s.windows.coef[0][0];
  const int n = nf * dim;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int fl = i / dim, d = i - fl * dim;
    const int64_t o = row_of[fl] + d;
    double wum = 0.0, row = 0.0;
    if (coef != 0.0) {
      const double wu = coef * with_ivar(s.vari[o]);
      wum += wu * s.mean[o];
      row += wu * coef;
    }
    s.par[f0 * dim + i] = wum / row;
  }
}

// ---------------------------------------------------------------- GV adjustment, one CTA per system
// MlpgGlobalVariance::{calc_gv, conv_gv, calc_hmmobj_derivative, next_step, parmgen}
// (mlpg.rs:168-292) on the solution left in the `g` plane by k_mlpg_solve, using the un-factorised
// band and W'U^-1 mu it stored. All passes are data-parallel over frames; the three sums per
// iteration are fixed-shape block reductions (deterministic; they differ from the reference's
// left-to-right sums only in rounding, far inside the 1e-9 parameter tolerance). The system lives
// in shared memory when it fits (smem_frames), else the same code runs on the global planes.
__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();  // red[] may still be read from the previous reduction
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  double t = 0.0;
  for (int i = 0; i < nw; ++i) t += red[i];
  return t;
}

// Three sums in one pass over the block: one barrier per call (the staging area alternates between two halves,
// so the values of the call before may still be read while this one writes).
__device__ __forceinline__ void block_sum3(double& a, double& b, double& c, double* red /* [2][3][8] */, int& phase) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
    c += __shfl_xor_sync(0xffffffffu, c, o);
  }
  const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  double* r = red + phase * 24;
  phase ^= 1;
  if ((threadIdx.x & 31) == 0) {
    r[w] = a;
    r[8 + w] = b;
    r[16 + w] = c;
  }
  __syncthreads();
  a = 0.0; b = 0.0; c = 0.0;
  for (int i = 0; i < nw; ++i) {
    a += r[i];
    b += r[8 + i];
    c += r[16 + i];
  }
}

template <int BW, bool sysmajor>
__global__ void __launch_bounds__(256, 3) k_mlpg_gv(BatchLayout b, StreamDev s, int64_t plane, int smem_frames) {
  extern __shared__ double sm[];
  __shared__ double red[8];
  __shared__ double red3[48];
  int phase = 0;
  const int sys = blockIdx.x;
  const int u = sys / s.dim, d = sys - u * s.dim;
  const int64_t soff = b.state_off[u], foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  const int dim = s.dim, tid = threadIdx.x, nt = blockDim.x;
  const bool msd = s.is_msd != 0;
  const int Tp = msd ? s.nvoiced[u] : T;
  const int32_t* const fstate = b.frame_state + foff;
  const int32_t* const vidx = msd ? s.vidx + foff : nullptr;
  double* out = s.par + foff * dim + d;
  if (msd)
    for (int t = tid; t < T; t += nt)
      if (s.left[foff + t] < 0) out[(int64_t)t * dim] = kNoData;  // Mask::fill (mask.rs:34-49)
  if (Tp == 0) return;

  // sysmajor (split path): the planes of this system are contiguous, see the plane map above
  double* const gbase = sysmajor ? s.scratch + foff * dim + (int64_t)d * T : s.scratch + foff * dim + d;
  const int gst = sysmajor ? 1 : dim;
  double* const g_par = gbase + (int64_t)(sysmajor ? 2 * BW + 2 : BW + 1) * plane;
  double* const g_wum = gbase + (int64_t)(sysmajor ? BW + 1 : BW) * plane;
  const bool in_smem = Tp <= smem_frames;
  double *PAR, *WUM, *GRAD, *O[BW];
  const int st = (sysmajor || in_smem) ? 1 : dim;  // element stride (compile-time 1 on the split path)
  uint8_t* SW = nullptr;
  if (in_smem) {
    PAR = sm;
#pragma unroll
    for (int j = 0; j < BW; ++j) O[j] = sm + (int64_t)(1 + j) * Tp;
    WUM = sm + (int64_t)(1 + BW) * Tp;
    GRAD = sm + (int64_t)(2 + BW) * Tp;
    SW = reinterpret_cast<uint8_t*>(sm + (int64_t)(3 + BW) * Tp);
    for (int k = tid; k < Tp; k += nt) {
      PAR[k] = g_par[(int64_t)k * gst];
      WUM[k] = g_wum[(int64_t)k * gst];
#pragma unroll
      for (int j = 0; j < BW; ++j) O[j][k] = gbase[(int64_t)(BW + 2 + j) * plane + (int64_t)k * gst];
      SW[k] = s.gv_switch[soff + fstate[msd ? vidx[k] : k]];
    }
    __syncthreads();
  } else {
    PAR = g_par;
    WUM = g_wum;
#pragma unroll
    for (int j = 0; j < BW; ++j) O[j] = gbase + (int64_t)(BW + 2 + j) * plane;
    GRAD = gbase + (int64_t)(sysmajor ? 2 * BW + 3 : 2 * BW + 2) * plane;
  }
  auto sw = [&](int k) -> bool { return in_smem ? SW[k] != 0 : s.gv_switch[soff + fstate[msd ? vidx[k] : k]] != 0; };

  double cnt = 0.0;
  for (int k = tid; k < Tp; k += nt) cnt += sw(k) ? 1.0 : 0.0;
  const int gv_length = (int)block_sum(cnt, red);
  if (gv_length > 0) {
    const double gv_mean = s.gv_mean[(int64_t)u * dim + d] * s.gv_weight;  // mlpg.rs:137-138
    const double gv_vari = s.gv_vari[(int64_t)u * dim + d];
    const double n = (double)gv_length;
    auto calc_gv = [&](double& mean, double& vari) {  // mlpg.rs:168-188
      double a = 0.0;
      for (int k = tid; k < Tp; k += nt)
        if (sw(k)) a += PAR[k * st];
      mean = block_sum(a, red) / n;
      a = 0.0;
      for (int k = tid; k < Tp; k += nt)
        if (sw(k)) {
          const double p = PAR[k * st];
          a += (p - mean) * (p - mean);
        }
      vari = block_sum(a, red) / n;
    };
    double m0;  // shift of the one-pass variance below: the mean (conv_gv keeps it, to rounding)
    {  // conv_gv (mlpg.rs:191-199)
      double mean, vari;
      calc_gv(mean, vari);
      m0 = mean;
      const double ratio = sqrt(gv_mean / vari);
      for (int k = tid; k < Tp; k += nt)
        if (sw(k)) PAR[k * st] = ratio * (PAR[k * st] - mean) + mean;
      __syncthreads();
    }
    const double wnorm = 1.0 / (double)((int64_t)s.nwin * (int64_t)Tp);
    const double len = (double)Tp;
    const double len2 = (double)((int64_t)Tp * (int64_t)Tp);
    const double lenm1 = (double)(Tp - 1);
    double step = 0.1, prevobj = 0.0;
    for (int it = 1; it <= 5; ++it) {
      // One pass, one barrier: the sums of calc_gv (mlpg.rs:168-188; variance about the shift m0, which is the
      // mean to rounding, so nothing cancels) and of calc_hmmobj_derivative (mlpg.rs:200-226) together.
      double a1 = 0.0, a2 = 0.0, h = 0.0;
      for (int k = tid; k < Tp; k += nt) {
        const double p = PAR[k * st];
        if (sw(k)) {
          a1 += p - m0;
          a2 += (p - m0) * (p - m0);
        }
        double g = O[0][k * st] * p;
#pragma unroll
        for (int i = 1; i < BW; ++i) {
          if (k + i < Tp) g += O[i][k * st] * PAR[(k + i) * st];
          if (k + 1 > i) g += O[i][(k - i) * st] * PAR[(k - i) * st];
        }
        GRAD[k * st] = g;
        h += 1.0 * wnorm * p * (WUM[k * st] - 0.5 * g);
      }
      block_sum3(a1, a2, h, red3, phase);
      const double dm = a1 / n;
      const double mean = m0 + dm, vari = a2 / n - dm * dm;
      m0 = mean;
      const double gvobj = -0.5 * 1.0 * vari * gv_vari * (vari - 2.0 * gv_mean);
      const double hmmobj = h;
      const double obj = -(hmmobj + gvobj);
      if (it > 1) {
        if (obj > prevobj)
          step *= 0.5;
        else if (obj < prevobj)
          step *= 1.2;
      }
      const double dv = -2.0 * gv_vari * (vari - gv_mean) / len;  // next_step (mlpg.rs:227-256)
      for (int k = tid; k < Tp; k += nt) {
        const double p = PAR[k * st];
        const double hh = -1.0 * wnorm * O[0][k * st] -
                          1.0 * 2.0 / len2 * (lenm1 * gv_vari * (vari - gv_mean) + 2.0 * gv_vari * (p - mean) * (p - mean));
        const double gk = GRAD[k * st], wk = WUM[k * st];
        double next_g;
        if (sw(k))
          next_g = 1.0 / hh * (1.0 * wnorm * (-gk + wk) + 1.0 * dv * (p - mean));
        else
          next_g = 1.0 / hh * (1.0 * wnorm * (-gk + wk));
        PAR[k * st] = p + step * next_g;
      }
      __syncthreads();
      prevobj = obj;
    }
  }
  for (int k = tid; k < Tp; k += nt) out[(int64_t)(msd ? vidx[k] : k) * dim] = PAR[k * st];
}

// The same adjustment in straight-line code, for systems of up to 128 E frames (split path, band width 3).
// k_mlpg_gv is bound by how many systems an SM holds (58 KB of shared memory each: three) and by what a thread does
// between two of its 17 block barriers (four elements, wrapped in index arithmetic and branches around the optional
// terms: 235 warp instructions per element and Newton step; profiles/r2_small_kernels_before_full_summary.txt). Here a
// system gets four warps, every thread owns E elements (tid + 128 j) for good and keeps their diagonal, W'U^-1 mu,
// gradient and switch in registers; only what a neighbour reads - the track and the two off-diagonals - lives in
// shared memory (31 KB), with a zero halo so that the terms the reference skips at the ends of the system become
// "+ 0 * 0" (an exact no-op) instead of a branch; a switched-off frame adds 0.0 to the GV sums the same way.
// Measured on the bench batch (35840 systems of ~1000 frames): 2.07 -> 1.58 ms with 0.74 G instructions instead of
// 1.34 G (1.82 ms at three CTAs per SM); eight warps x five elements 2.06 ms, two x twenty 2.32 ms, and with all six arrays in shared memory
// (58 KB per system again) 6.1 / 3.2 / 2.2 / 1.9 ms for one / two / four / eight warps: the kernel lives on the
// number of warps an SM can hold, registers and shared memory both count. Arithmetic per element as in k_mlpg_gv;
// the sums of a Newton step are taken over other groupings of the elements, so they differ from that kernel's (and
// from the reference's left-to-right sums) in rounding only.
// Six CTAs per SM (80 registers, ~140 bytes of spills outside the element loops): measured 2.54 / 1.82 / 1.81 /
// 1.74 / 1.58 / 1.57 ms for at least 1 / 3 / 4 / 5 / 6 / 7 CTAs per SM - occupancy is worth more than the spills.
template <int NW, int E>
__global__ void __launch_bounds__(32 * NW, 6) k_mlpg_gv_lin(BatchLayout b, StreamDev s, int64_t plane) {
  constexpr int BW = 3, H = 2, NT = 32 * NW, N = NT * E + 2 * H;
  __shared__ double Ps[N], O1s[N], O2s[N];
  __shared__ double red3[48];
  int phase = 0;
  double* const P = Ps + H;
  double* const O1 = O1s + H;
  double* const O2 = O2s + H;
  const int sys = blockIdx.x, tid = threadIdx.x;
  const int u = sys / s.dim, d = sys - u * s.dim;
  const int64_t soff = b.state_off[u], foff = b.frame_off[u];
  const int T = (int)(b.frame_off[u + 1] - foff);
  const int dim = s.dim;
  const bool msd = s.is_msd != 0;
  const int Tp = msd ? s.nvoiced[u] : T;
  const int32_t* const fstate = b.frame_state + foff;
  const int32_t* const vidx = msd ? s.vidx + foff : nullptr;
  double* out = s.par + foff * dim + d;
  if (msd)
    for (int t = tid; t < T; t += NT)
      if (s.left[foff + t] < 0) out[(int64_t)t * dim] = kNoData;  // Mask::fill (mask.rs:34-49)
  if (Tp == 0) return;
  const double* const gbase = s.scratch + foff * dim + (int64_t)d * T;  // system-major planes of this system
  double o0[E], wum[E], gr[E];
  unsigned swm = 0;
  if (tid < H) {
    Ps[tid] = 0.0; O1s[tid] = 0.0; O2s[tid] = 0.0;
    Ps[N - 1 - tid] = 0.0; O1s[N - 1 - tid] = 0.0; O2s[N - 1 - tid] = 0.0;
  }
#pragma unroll
  for (int j = 0; j < E; ++j) {
    const int k = tid + j * NT;
    const bool in = k < Tp;
    P[k] = in ? gbase[(int64_t)(2 * BW + 2) * plane + k] : 0.0;
    wum[j] = in ? gbase[(int64_t)(BW + 1) * plane + k] : 0.0;
    o0[j] = in ? gbase[(int64_t)(BW + 2) * plane + k] : 1.0;  // a pivot of 1 keeps the padding finite
    O1[k] = in ? gbase[(int64_t)(BW + 3) * plane + k] : 0.0;  // the builder left 0 where k + i >= Tp
    O2[k] = in ? gbase[(int64_t)(BW + 4) * plane + k] : 0.0;
    if (in && s.gv_switch[soff + fstate[msd ? vidx[k] : k]] != 0) swm |= 1u << j;
    gr[j] = 0.0;
  }
  __syncthreads();
  double cnt = 0.0, a = 0.0, dummy = 0.0;
#pragma unroll
  for (int j = 0; j < E; ++j) {
    const bool on = swm >> j & 1u;
    cnt += on ? 1.0 : 0.0;
    a += on ? P[tid + j * NT] : 0.0;
  }
  block_sum3(cnt, a, dummy, red3, phase);
  const int gv_length = (int)cnt;
  if (gv_length > 0) {
    const double gv_mean = s.gv_mean[(int64_t)u * dim + d] * s.gv_weight;  // mlpg.rs:137-138
    const double gv_vari = s.gv_vari[(int64_t)u * dim + d];
    const double n = (double)gv_length;
    double m0;
    {  // calc_gv (mlpg.rs:168-188) + conv_gv (:191-199)
      const double mean = a / n;
      double v = 0.0, d1 = 0.0, d2 = 0.0;
#pragma unroll
      for (int j = 0; j < E; ++j) {
        const double pk = P[tid + j * NT];
        v += (swm >> j & 1u) ? (pk - mean) * (pk - mean) : 0.0;
      }
      block_sum3(v, d1, d2, red3, phase);
      const double vari = v / n;
      m0 = mean;
      const double ratio = sqrt(gv_mean / vari);
#pragma unroll
      for (int j = 0; j < E; ++j)
        if (swm >> j & 1u) P[tid + j * NT] = ratio * (P[tid + j * NT] - mean) + mean;
      __syncthreads();
    }
    const double wnorm = 1.0 / (double)((int64_t)s.nwin * (int64_t)Tp);
    const double len = (double)Tp;
    const double len2 = (double)((int64_t)Tp * (int64_t)Tp);
    const double lenm1 = (double)(Tp - 1);
    double step = 0.1, prevobj = 0.0;
    for (int it = 1; it <= 5; ++it) {
      double a1 = 0.0, a2 = 0.0, h = 0.0;
#pragma unroll
      for (int j = 0; j < E; ++j) {
        const int k = tid + j * NT;
        const double pk = P[k];
        const double dp = pk - m0;
        const bool on = swm >> j & 1u;
        a1 += on ? dp : 0.0;
        a2 += on ? dp * dp : 0.0;
        double g = o0[j] * pk;  // calc_hmmobj_derivative (mlpg.rs:200-226), same order of terms
        g += O1[k] * P[k + 1];
        g += O1[k - 1] * P[k - 1];
        g += O2[k] * P[k + 2];
        g += O2[k - 2] * P[k - 2];
        gr[j] = g;
        h += 1.0 * wnorm * pk * (wum[j] - 0.5 * g);
      }
      block_sum3(a1, a2, h, red3, phase);  // its barrier also separates the reads of P above from the writes below
      const double dm = a1 / n;
      const double mean = m0 + dm, vari = a2 / n - dm * dm;
      m0 = mean;
      const double gvobj = -0.5 * 1.0 * vari * gv_vari * (vari - 2.0 * gv_mean);
      const double obj = -(h + gvobj);
      if (it > 1) {
        if (obj > prevobj)
          step *= 0.5;
        else if (obj < prevobj)
          step *= 1.2;
      }
      const double dv = -2.0 * gv_vari * (vari - gv_mean) / len;  // next_step (mlpg.rs:227-256)
      const double c2 = 1.0 * 2.0 / len2, c3 = lenm1 * gv_vari * (vari - gv_mean);
#pragma unroll
      for (int j = 0; j < E; ++j) {
        const int k = tid + j * NT;
        const double pk = P[k];
        const double hh = -1.0 * wnorm * o0[j] - c2 * (c3 + 2.0 * gv_vari * (pk - mean) * (pk - mean));
        const double base = 1.0 * wnorm * (-gr[j] + wum[j]);
        const double next_g = 1.0 / hh * ((swm >> j & 1u) ? base + 1.0 * dv * (pk - mean) : base);
        if (k < Tp) P[k] = pk + step * next_g;
      }
      __syncthreads();
      prevobj = obj;
    }
  }
#pragma unroll
  for (int j = 0; j < E; ++j) {
    const int k = tid + j * NT;
    if (k < Tp) out[(int64_t)(msd ? vidx[k] : k) * dim] = P[k];
  }
}

}  // namespace

namespace {
__global__ void __launch_bounds__(256) k_gather_states(GatherStream g, int64_t total) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= total) return;
  const int64_t i = e / g.width;
  const int k = (int)(e - i * g.width);
  const int32_t* r = g.rows + i * g.nvoices;
  int64_t row = r[0];
  double m = g.tab_mean[row * g.width + k] * g.w[0];
  double v = g.tab_vari[row * g.width + k] * g.w[0];
  double ms = (g.is_msd && k == 0) ? g.w[0] * g.tab_msd[row] : 0.0;
  for (int q = 1; q < g.nvoices; ++q) {
    row = r[q];
    m += g.w[q] * g.tab_mean[row * g.width + k];  // -fmad=false: multiply, then add
    v += g.w[q] * g.tab_vari[row * g.width + k];
    if (g.is_msd && k == 0) ms += g.w[q] * g.tab_msd[row];
  }
  if (g.half_tone && k == 0) {
    m = m + g.half_tone_add;
    m = m < g.lf0_min ? g.lf0_min : (m > g.lf0_max ? g.lf0_max : m);
  }
  g.mean[e] = m;
  g.vari[e] = v;
  if (g.is_msd && k == 0) g.msd[i] = ms;
}
}  // namespace

int launch_gather_states(const GatherStream& g, int64_t total_states, cudaStream_t stream) {
  const int64_t total = total_states * g.width;
  if (total <= 0) return 0;
  if (g.nvoices < 1 || g.nvoices > kMaxVoices) return (int)cudaErrorInvalidValue;
  k_gather_states<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(g, total);
  return (int)cudaGetLastError();
}

int launch_expand(const BatchLayout& b, const StreamSet& set, cudaStream_t stream) {
  if (b.n_utt == 0) return 0;
  k_expand<<<b.n_utt, kExpandThreads, 0, stream>>>(b, set);
  return (int)cudaGetLastError();
}

int launch_mlpg(const BatchLayout& b, const StreamDev& s, int64_t total_frames, int64_t total_states,
                cudaStream_t stream) {
  if (b.n_utt == 0 || s.dim == 0) return 0;
  const int64_t plane = total_frames * s.dim;
  const int64_t nthreads = (int64_t)b.n_utt * s.dim;
  const int threads = 128;
  const int blocks = (int)((nthreads + threads - 1) / threads);
  const WindowSet& w = s.windows;
  const bool shape1 = w.nwin == 1 && w.width[0] == 1;
  const bool shape133 = w.nwin == 3 && w.width[0] == 1 && w.width[1] == 3 && w.width[2] == 3;
  if (shape1 && !s.is_msd && !s.has_gv) {  // every frame is its own system
    if (total_frames == 0 || s.dim == 0) return 0;
    k_mlpg_static<<<(unsigned)((total_frames + kStaticFrames - 1) / kStaticFrames), 256, 0, stream>>>(b, s, total_frames);
    return (int)cudaGetLastError();
  }
  if (shape1 || shape133) {
    if (shape1) {
      k_mlpg_solve<kShape1><<<blocks, threads, 0, stream>>>(b, s, plane);
    } else {
      if (s.ivar == nullptr) return (int)cudaErrorInvalidValue;
      const int64_t nv = total_states * s.dim * s.nwin;
      if (nv > 0) k_state_ivar<<<(unsigned)((nv + 255) / 256), 256, 0, stream>>>(s.vari, s.ivar, nv);
      if (b.max_frames > 0)
        k_band_build<<<dim3((unsigned)((b.max_frames + 127) / 128), (unsigned)b.n_utt), 128, 0, stream>>>(b, s, plane);
      k_ldl<<<blocks, threads, 0, stream>>>(b, s, plane);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess || !s.has_gv) return (int)e;
    // GV: one CTA per system, the system in shared memory when it fits
    const int bw = shape1 ? 1 : 3;
    const size_t per_frame = (size_t)(3 + bw) * sizeof(double) + 1;
    const size_t cap = 200 * 1024;
    int smem_frames = (int)((cap - 16) / per_frame);
    if (smem_frames > b.max_frames) smem_frames = b.max_frames;
    const size_t smem = (size_t)smem_frames * per_frame + 16;
    // per device and cheap, so set on every launch
    cudaFuncSetAttribute(k_mlpg_gv<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cap);
    cudaFuncSetAttribute(k_mlpg_gv<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cap);
    const unsigned g = (unsigned)nthreads;
    if (bw == 1) k_mlpg_gv<1, false><<<g, 256, smem, stream>>>(b, s, plane, smem_frames);
    else if (b.max_frames <= 128 * 8) k_mlpg_gv_lin<4, 8><<<g, 128, 0, stream>>>(b, s, plane);
    else if (b.max_frames <= 128 * 10) k_mlpg_gv_lin<4, 10><<<g, 128, 0, stream>>>(b, s, plane);
    else if (b.max_frames <= 128 * 12) k_mlpg_gv_lin<4, 12><<<g, 128, 0, stream>>>(b, s, plane);
    else k_mlpg_gv<3, true><<<g, 256, smem, stream>>>(b, s, plane, smem_frames);
    return (int)cudaGetLastError();
  }
  switch (s.windows.band_width()) {  // other window shapes: the general kernel
    case 1: k_mlpg<1><<<blocks, threads, 0, stream>>>(b, s, plane); break;
    case 3: k_mlpg<3><<<blocks, threads, 0, stream>>>(b, s, plane); break;
    case 5: k_mlpg<5><<<blocks, threads, 0, stream>>>(b, s, plane); break;
    default: return (int)cudaErrorInvalidValue;
  }
  return (int)cudaGetLastError();
}

}  // namespace jbk
