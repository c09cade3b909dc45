// MLPG solve of one non-MSD stream without any allocation: the band planes that k_band_build / k_ldl (mlpg.cu)
// keep in the engine's cudaMalloc'ed work arena live here in a file-scope __device__ array of fixed size.
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). Declaration and contract:
// include/jbonsai_b200.h (jb_mlpg_solve_noalloc).
//
// Arithmetic: MlpgMatrix::calc_wuw_and_wum, ldl_factorization, substitutions (src/mlpg_adjust/mlpg.rs:25-115) in the
// reference's operation order, MeanVari::with_ivar (src/model/mean_vari.rs:21-31) and the boundary rule of
// MlpgAdjust::create (src/mlpg_adjust/mod.rs:66-76) for a stream whose every frame is kept. Same results as
// jb_engine_mlpg_create on the same input (tests/test_zz_no_alloc.py).
//
// The scratch array is shared by every call on a device and keeps the last call's contents. Nothing in it has to
// start at zero: k_na_band writes every cell (band planes 0..band-1 and the W'U^-1 mu plane, n_frames x dim each) that
// k_na_ldl reads in the same call, on the same stream. Two calls whose kernels overlap on one GPU (different
// streams, or different host threads) race on it.
#include <mutex>

#include "kernels.h"

namespace jbk {

namespace {

constexpr int kNaPlanes = kMaxWinWidth + 1;  // band (<= 5 entries per row) + the right-hand side / forward solution
constexpr size_t kNaScratchDoubles = (size_t)kNaPlanes * kMlpgNoAllocMaxCells;  // 6 MiB

// In host code this name is only a handle for cudaGetSymbolAddress: it is NOT a device address.
static __device__ __align__(256) double g_na_scratch[kNaScratchDoubles];

constexpr int kNaMaxDevices = 64;
std::mutex g_na_mutex;
double* g_na_ptr[kNaMaxDevices] = {};  // device address of g_na_scratch on every device set up so far

__device__ __forceinline__ double na_with_ivar(double vari) {  // mean_vari.rs:21-31
  if (fabs(vari) > 1e19) return 0.0;
  if (fabs(vari) < 1e-19) return 1e38;
  return 1.0 / vari;
}

// One thread per (frame, dimension): row t of W'U^-1 W (upper band) and of W'U^-1 mu. mlpg.rs:25-71.
__global__ void __launch_bounds__(256) k_na_band(const double* __restrict__ mean, const double* __restrict__ vari,
                                                 double* __restrict__ planes, WindowSet ws, int band, long long T,
                                                 int dim) {
  const long long cells = T * dim;
  const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= cells) return;
  const long long t = cell / dim;
  const int d = (int)(cell - t * dim);
  const long long L = (long long)ws.nwin * dim;
  double wum = 0.0;
  double wuw[kMaxWinWidth];
#pragma unroll
  for (int k = 0; k < kMaxWinWidth; ++k) wuw[k] = 0.0;
#pragma unroll
  for (int i = 0; i < kMaxWin; ++i) {
    if (i >= ws.nwin) break;
    const int ww = ws.width[i], half = ww / 2;
    const int lw = half, rw = ww - half - 1;
#pragma unroll
    for (int index = kMaxWinWidth - 1; index >= 0; --index) {  // window.iter_rev(0)
      if (index >= ww) continue;
      const double coef = ws.coef[i][index];
      if (coef == 0.0) continue;
      const long long idx = t - (index - half);
      if (idx < 0 || idx >= T) continue;
      double ivar = na_with_ivar(vari[idx * L + (long long)dim * i + d]);
      if (i != 0 && (idx < lw || T - 1 - idx < rw)) ivar = 0.0;  // mod.rs:66-76 with no masked frame
      const double wu = coef * ivar;
      wum += wu * mean[idx * L + (long long)dim * i + d];
      bool stop = false;  // the reference leaves the inner loop at the first entry past the end (mlpg.rs:62-64)
#pragma unroll
      for (int inner = kMaxWinWidth - 1; inner >= 0; --inner) {  // window.iter_rev(index)
        if (stop || inner >= ww || inner < index) continue;
        const double c2 = ws.coef[i][inner];
        if (c2 == 0.0) continue;
        const int j = inner - index;
        if (t + j >= T) {
          stop = true;
          continue;
        }
        wuw[j] += wu * c2;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < kMaxWinWidth; ++k)
    if (k < band) planes[(long long)k * cells + cell] = wuw[k];
  planes[(long long)band * cells + cell] = wum;
}

// One thread per dimension, serial along time: LDL factorisation in place, forward substitution into the
// right-hand-side plane, backward substitution into `out`. mlpg.rs:79-115.
__global__ void __launch_bounds__(128) k_na_ldl(double* __restrict__ planes, double* __restrict__ out, int band,
                                                long long T, int dim) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= dim) return;
  const long long cells = T * dim;
  double* const g = planes + (long long)band * cells;
#define NA_P(k, tt) planes[(long long)(k) * cells + (tt) * dim + d]
  for (long long t = 0; t < T; ++t) {
    double w0 = NA_P(0, t);
    for (int i = 1; i < band && i <= t; ++i) w0 -= NA_P(i, t - i) * NA_P(i, t - i) * NA_P(0, t - i);
    NA_P(0, t) = w0;
    for (int i = 1; i < band; ++i) {
      double wi = NA_P(i, t);
      for (int j = 1; j < band - i && j <= t; ++j) wi -= NA_P(j, t - j) * NA_P(i + j, t - j) * NA_P(0, t - j);
      NA_P(i, t) = wi / w0;
    }
    double gt = g[t * dim + d];
    for (int i = 1; i < band && i <= t; ++i) gt -= NA_P(i, t - i) * g[(t - i) * dim + d];
    g[t * dim + d] = gt;
  }
  for (long long t = T - 1; t >= 0; --t) {
    double p = g[t * dim + d] / NA_P(0, t);
    for (int i = 1; i < band && i < T - t; ++i) p -= NA_P(i, t) * out[(t + i) * dim + d];
    out[t * dim + d] = p;
  }
#undef NA_P
}

}  // namespace

// Host side of jb_mlpg_solve_noalloc (capi.cc validates the plain arguments and keeps the error text). Returns a
// cudaError_t; *what names the step that failed. Calls no cudaMalloc / cudaFree / allocating routine.
int mlpg_solve_noalloc(int device, const double* d_mean, const double* d_vari, long long n_frames, int dim,
                       const WindowSet& ws, double* d_out, cudaStream_t stream, const char** what) {
  *what = "";
  if (device < 0 || device >= kNaMaxDevices) {
    *what = "device ordinal out of range";
    return (int)cudaErrorInvalidDevice;
  }
  int prev = -1;
  cudaError_t e = cudaGetDevice(&prev);
  if (e != cudaSuccess) {
    *what = "cudaGetDevice";
    return (int)e;
  }
  if (prev != device && (e = cudaSetDevice(device)) != cudaSuccess) {
    *what = "cudaSetDevice";
    return (int)e;
  }
  double* scratch = nullptr;
  {
    // once per device, with that device current: the device address of the __device__ array
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (g_na_ptr[device] == nullptr) {
      double* p = nullptr;
      e = cudaGetSymbolAddress((void**)&p, g_na_scratch);
      if (e == cudaSuccess && p == nullptr) e = cudaErrorInvalidSymbol;
      if (e == cudaSuccess) g_na_ptr[device] = p;
    }
    scratch = g_na_ptr[device];
  }
  if (e != cudaSuccess) {
    *what = "cudaGetSymbolAddress(scratch)";
  } else if (n_frames > 0) {
    const int band = ws.band_width();
    const long long cells = n_frames * dim;
    k_na_band<<<(unsigned)((cells + 255) / 256), 256, 0, stream>>>(d_mean, d_vari, scratch, ws, band, n_frames, dim);
    e = cudaGetLastError();
    if (e != cudaSuccess) {
      *what = "launch of k_na_band";
    } else {
      k_na_ldl<<<(unsigned)((dim + 127) / 128), 128, 0, stream>>>(scratch, d_out, band, n_frames, dim);
      e = cudaGetLastError();
      if (e != cudaSuccess) *what = "launch of k_na_ldl";
    }
  }
  if (prev != device && prev >= 0) cudaSetDevice(prev);
  return (int)e;
}

}  // namespace jbk
