// Engine implementation: host batching layer + CUDA pipeline driver (no CPU fallback).
#include "engine.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <thread>

#include <cuda_runtime.h>

#include "cuda/kernels.h"

namespace jb {

#define JB_CUDA(x)                                                                          \
  do {                                                                                      \
    cudaError_t e_ = (x);                                                                   \
    if (e_ != cudaSuccess)                                                                  \
      throw Error(kErrCuda, std::string("CUDA error: ") + cudaGetErrorString(e_) + " in " #x); \
  } while (0)
#define JB_LAUNCH(x)                                                                        \
  do {                                                                                      \
    int e_ = (x);                                                                           \
    if (e_ != 0)                                                                            \
      throw Error(kErrCuda, std::string("CUDA launch error: ") +                            \
                                cudaGetErrorString((cudaError_t)e_) + " in " #x);           \
  } while (0)

namespace {

// resident warps per scheduler of k_mlsa_v3 (255 registers: two fit)
constexpr int kSegWarpsPerSmsp = 2;
constexpr int kRepairWarmFactor = 2;
constexpr double kMglsaCutRelTolerance = 1e-9;
constexpr int kMglsaWarmFactor = 2;  // stage != 0: the all-pole MGLSA filter of an LSP frame forgets more slowly than the
                                     // MLSA filter (synthetic tracks: 40 frames leave 1e0, 80 frames 1e-12 at the cuts)
constexpr int kSegMinFrames = 48;    // shortest own range worth a segment of its own

double now_ms() {
  using namespace std::chrono;
  return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

// ---- pinned host pool (process-wide): outputs and staging blobs are recycled, because pinning
// gigabytes per call would cost more than the kernels.
struct PinnedPool {
  std::mutex mu;
  std::map<void*, size_t> live;                  // handed out
  std::multimap<size_t, void*> cached;           // free, by capacity
  size_t cached_bytes = 0;
  // Free blocks kept for reuse. One f64 batch of BASELINE config 4 (16384 utterances in ONE call) is 31.5 GB:
  // pinning that afresh costs ~14 s per call, so the cap must sit above it (JB_PINNED_CACHE_GB overrides).
  const size_t kMaxCached = (size_t)(getenv("JB_PINNED_CACHE_GB") ? atoi(getenv("JB_PINNED_CACHE_GB")) : 48) << 30;
  void* alloc(size_t bytes) {
    if (bytes == 0) bytes = 8;
    {
      std::lock_guard<std::mutex> g(mu);
      auto it = cached.lower_bound(bytes);
      if (it != cached.end() && it->first <= bytes * 2 + (1 << 20)) {
        void* p = it->second;
        size_t cap = it->first;
        cached.erase(it);
        cached_bytes -= cap;
        live[p] = cap;
        return p;
      }
    }
    void* p = nullptr;
    cudaError_t e = cudaMallocHost(&p, bytes);
    if (e != cudaSuccess)
      throw Error(kErrCuda, std::string("cudaMallocHost failed: ") + cudaGetErrorString(e));
    std::lock_guard<std::mutex> g(mu);
    live[p] = bytes;
    return p;
  }
  void release(void* p) {
    if (!p) return;
    size_t cap = 0;
    {
      std::lock_guard<std::mutex> g(mu);
      auto it = live.find(p);
      if (it == live.end()) return;
      cap = it->second;
      live.erase(it);
      if (cached_bytes + cap <= kMaxCached) {
        cached.emplace(cap, p);
        cached_bytes += cap;
        return;
      }
    }
    cudaFreeHost(p);
  }
};
PinnedPool& pinned() {
  static PinnedPool* p = new PinnedPool();  // leaked on purpose: outlives static destructors
  return *p;
}
size_t pinned_trim() {  // frees every cached (not handed out) pinned block; returns the bytes released
  PinnedPool& pp = pinned();
  std::multimap<size_t, void*> drop;
  size_t bytes = 0;
  {
    std::lock_guard<std::mutex> g(pp.mu);
    drop.swap(pp.cached);
    bytes = pp.cached_bytes;
    pp.cached_bytes = 0;
  }
  for (auto& kv : drop) cudaFreeHost(kv.second);
  return bytes;
}

size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Host threads the front-end and the batching layer may use: all cores by default; a process that shares the
// box with others (one rank per GPU under torchrun) caps it with jb_set_host_threads / JB_HOST_THREADS, so that
// eight ranks do not run eight times the core count.
std::atomic<int> g_host_threads{0};
size_t host_threads() {
  int n = g_host_threads.load();
  if (n <= 0) {
    static const int env = getenv("JB_HOST_THREADS") ? atoi(getenv("JB_HOST_THREADS")) : 0;
    n = env > 0 ? env : (int)std::max(1u, std::thread::hardware_concurrency());
  }
  return (size_t)n;
}

// A thread that shares the host with `t_host_share - 1` peers doing the same (one host thread per GPU of a
// multi-device synthesize_batch call) takes its share of the cores only.
thread_local size_t t_host_share = 1;

// Runs fn(0..n-1) on all host cores; rethrows the first jb::Error.
template <class F>
void parallel_for(size_t n, F fn) {
  size_t nthreads = std::min<size_t>(std::max<size_t>(1, host_threads() / t_host_share), n / 8 + 1);
  if (nthreads <= 1) {
    for (size_t i = 0; i < n; ++i) fn(i);
    return;
  }
  std::atomic<size_t> next{0};
  std::mutex emu;
  std::unique_ptr<Error> err;
  std::vector<std::thread> th;
  for (size_t t = 0; t < nthreads; ++t)
    th.emplace_back([&]() {
      for (;;) {
        const size_t i = next.fetch_add(1);
        if (i >= n) return;
        try {
          fn(i);
        } catch (const Error& e) {
          std::lock_guard<std::mutex> g(emu);
          if (!err) err.reset(new Error(e));
        } catch (const std::exception& e) {
          std::lock_guard<std::mutex> g(emu);
          if (!err) err.reset(new Error(kErrArgument, e.what()));
        }
      }
    });
  for (auto& t : th) t.join();
  if (err) throw *err;
}

struct Carver {  // carve aligned sections out of one allocation
  size_t off = 0;
  size_t take(size_t bytes) {
    size_t o = off;
    off = align_up(off + bytes);
    return o;
  }
};

}  // namespace

// ---------------------------------------------------------------- per-GPU context
class DeviceCtx {
 public:
  int device;
  cudaStream_t stream = nullptr;
  cudaStream_t stream2 = nullptr;  // side stream: LF0 solve -> pitch -> pulse train run beside MCP/LPF
  cudaEvent_t evx[2];
  cudaStream_t stream_copy = nullptr;  // device-to-host copies of finished filter waves
  cudaEvent_t evw[2];
  std::mutex mu;  // one in-flight batch per (engine, device)
  double* d_noise = nullptr;
  size_t noise_len = 0;
  // Timing events and the pinned cut-check word come in kMaxChunks slots: synthesize_batch keeps
  // several sub-batches in flight on one device and reads each one's numbers at the end.
  static constexpr int kMaxChunks = 4;
  cudaEvent_t evsets[kMaxChunks][8];
  cudaEvent_t* ev = evsets[0];
  void use_slot(int i) {
    ev = evsets[i];
    h_worst = h_worst_base + i;
    h_fail = h_fail_base + i;
  }
  std::multimap<size_t, void*> cached;  // freed device blocks
  int sm_count = 148;
  double* h_worst_base = nullptr;
  double* h_worst = nullptr;  // pinned: MLSA variant 3 cut-check result (of the current slot)
  int loud_strikes = 0;       // batches whose loudness-sized warm-up left many cuts to repair (2: rule off for this device)
  int* h_fail_base = nullptr;
  int* h_fail = nullptr;      // pinned: number of cuts that failed the check (of the current slot)
  // leaf pdf tables of the engine's voices, resident for the life of the context (device gather)
  double* d_tab_mean[3] = {nullptr, nullptr, nullptr};
  double* d_tab_vari[3] = {nullptr, nullptr, nullptr};
  double* d_tab_msd[3] = {nullptr, nullptr, nullptr};
  bool tables_loaded = false;
  void load_tables(const std::vector<StreamTable>& tabs) {
    if (tables_loaded) return;
    for (size_t s = 0; s < tabs.size() && s < 3; ++s) {
      const StreamTable& t = tabs[s];
      if (!t.regular || t.rows == 0) continue;
      JB_CUDA(cudaMalloc((void**)&d_tab_mean[s], t.mean.size() * 8));
      JB_CUDA(cudaMalloc((void**)&d_tab_vari[s], t.vari.size() * 8));
      JB_CUDA(cudaMemcpy(d_tab_mean[s], t.mean.data(), t.mean.size() * 8, cudaMemcpyHostToDevice));
      JB_CUDA(cudaMemcpy(d_tab_vari[s], t.vari.data(), t.vari.size() * 8, cudaMemcpyHostToDevice));
      if (t.is_msd) {
        JB_CUDA(cudaMalloc((void**)&d_tab_msd[s], t.msd.size() * 8));
        JB_CUDA(cudaMemcpy(d_tab_msd[s], t.msd.data(), t.msd.size() * 8, cudaMemcpyHostToDevice));
      }
    }
    tables_loaded = true;
  }

  explicit DeviceCtx(int dev) : device(dev) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
      throw Error(kErrCuda, std::string("no CUDA device available (") +
                                (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") +
                                "); jbonsai_b200 has no CPU fallback");
    if (dev < 0 || dev >= n) throw Error(kErrCuda, "CUDA device index out of range");
    JB_CUDA(cudaSetDevice(dev));
    JB_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    {
      // The LF0 -> pitch -> pulse chain is a string of small latency-bound kernels; at a higher
      // priority its blocks slip in between those of the chip-filling MCP kernels instead of
      // queueing behind them.
      int least = 0, greatest = 0;
      JB_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
      JB_CUDA(cudaStreamCreateWithPriority(&stream2, cudaStreamNonBlocking, greatest));
    }
    JB_CUDA(cudaStreamCreateWithFlags(&stream_copy, cudaStreamNonBlocking));
    for (auto& x : evx) JB_CUDA(cudaEventCreateWithFlags(&x, cudaEventDisableTiming));
    for (auto& x : evw) JB_CUDA(cudaEventCreateWithFlags(&x, cudaEventDisableTiming));
    for (auto& set : evsets)
      for (auto& x : set) JB_CUDA(cudaEventCreate(&x));
    JB_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    JB_CUDA(cudaMallocHost((void**)&h_worst_base, kMaxChunks * sizeof(double)));
    JB_CUDA(cudaMallocHost((void**)&h_fail_base, kMaxChunks * sizeof(int)));
    for (int i = 0; i < kMaxChunks; ++i) h_worst_base[i] = 0.0;
    for (int i = 0; i < kMaxChunks; ++i) h_fail_base[i] = 0;
    use_slot(0);
  }
  ~DeviceCtx() {
    cudaSetDevice(device);
    for (auto& kv : cached) cudaFree(kv.second);
    if (d_noise) cudaFree(d_noise);
    for (int s = 0; s < 3; ++s) {
      if (d_tab_mean[s]) cudaFree(d_tab_mean[s]);
      if (d_tab_vari[s]) cudaFree(d_tab_vari[s]);
      if (d_tab_msd[s]) cudaFree(d_tab_msd[s]);
    }
    if (h_worst_base) cudaFreeHost(h_worst_base);
    if (h_fail_base) cudaFreeHost(h_fail_base);
    for (auto& set : evsets)
      for (auto& x : set) cudaEventDestroy(x);
    for (auto& x : evx) cudaEventDestroy(x);
    for (auto& x : evw) cudaEventDestroy(x);
    if (stream_copy) cudaStreamDestroy(stream_copy);
    if (stream2) cudaStreamDestroy(stream2);
    if (stream) cudaStreamDestroy(stream);
  }
  void* dalloc(size_t bytes, size_t* cap) {
    if (bytes == 0) bytes = 256;
    auto it = cached.lower_bound(bytes);
    if (it != cached.end() && it->first <= bytes * 2 + (1 << 20)) {
      void* p = it->second;
      *cap = it->first;
      cached.erase(it);
      return p;
    }
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
      for (auto& kv : cached) cudaFree(kv.second);  // drop the cache and retry once
      cached.clear();
      JB_CUDA(cudaMalloc(&p, bytes));
    }
    *cap = bytes;
    return p;
  }
  void dfree(void* p, size_t cap) {
    if (p) cached.emplace(cap, p);
  }
};

// ---------------------------------------------------------------- engine basics
Engine::Engine(VoiceSet v) : voices_(std::move(v)) {
  condition.load_model(voices_);
  tables_ = build_stream_tables(voices_);
  for (auto& t : tables_)
    if (voices_.size() > (size_t)jbk::kMaxVoices) t.regular = false;
}
Engine::~Engine() = default;

std::unique_ptr<Engine> Engine::load(const std::vector<std::string>& paths) {
  std::vector<std::vector<uint8_t>> bufs;
  for (const auto& p : paths) bufs.push_back(read_file(p));
  return load_from_bytes(bufs);
}

std::unique_ptr<Engine> Engine::load_from_bytes(const std::vector<std::vector<uint8_t>>& bufs) {
  std::vector<std::shared_ptr<Voice>> voices;
  for (const auto& b : bufs) voices.push_back(parse_htsvoice(b.data(), b.size()));
  return std::unique_ptr<Engine>(new Engine(VoiceSet(std::move(voices))));
}

void Engine::set_devices(const std::vector<int>& devices) {
  if (devices.empty()) throw Error(kErrArgument, "device list is empty");
  devices_ = devices;
}

DeviceCtx* Engine::ctx(int device) const {
  std::lock_guard<std::mutex> g(ctx_mu_);
  auto it = ctx_.find(device);
  if (it == ctx_.end()) it = ctx_.emplace(device, std::unique_ptr<DeviceCtx>(new DeviceCtx(device))).first;
  return it->second.get();
}

void Engine::free_output(void* p) { pinned().release(p); }
void* Engine::alloc_output(size_t bytes) { return pinned().alloc(bytes ? bytes : 8); }
void Engine::set_host_threads(int n) { g_host_threads.store(n > 0 ? n : 0); }
size_t Engine::trim_pools() { return pinned_trim(); }

// Random::{nrandom, fill_queue} + rnd (src/vocoder/excitation.rs:197-237), run on the host so
// that ln/sqrt come from the same libm the reference uses; extended lazily, 64 values a time.
const std::vector<double>& Engine::noise_table(size_t n) const {  // caller holds noise_mu_
  auto rnd = [&]() {
    rng_next_ = rng_next_ * 1103515245ull + 12345ull;
    uint64_t r = (rng_next_ / 65536ull) % 32768ull;
    return (double)r / 32767.0;
  };
  if (noise_.capacity() < n) noise_.reserve(n + n / 4 + 64);
  while (noise_.size() < n) {
    double q[64], s[32];
    int i = 0;
    while (i < 32) {
      double r1 = 2.0 * rnd() - 1.0;
      double r2 = 2.0 * rnd() - 1.0;
      double ss = r1 * r1 + r2 * r2;
      if (0.0 < ss && ss < 1.0) {
        q[2 * i] = r1;
        q[2 * i + 1] = r2;
        s[i] = ss;
        ++i;
      }
    }
    for (int k = 0; k < 32; ++k) {
      double m = std::sqrt(-2.0 * std::log(s[k]) / s[k]);
      q[2 * k] *= m;
      q[2 * k + 1] *= m;
    }
    noise_.insert(noise_.end(), q, q + 64);
  }
  return noise_;
}

std::vector<double> Engine::noise_prefix(size_t n) const {
  std::lock_guard<std::mutex> g(noise_mu_);
  const std::vector<double>& t = noise_table(n);
  return std::vector<double>(t.begin(), t.begin() + n);
}

std::vector<UtterancePack> Engine::build_packs(const std::vector<std::vector<std::string>>& all, size_t first,
                                               size_t count) const {
  if (first > all.size()) first = all.size();
  if (count > all.size() - first) count = all.size() - first;
  struct Range {  // utterances [first, first + count) of `all`, indexed from 0
    const std::vector<std::string>* p;
    size_t n;
    size_t size() const { return n; }
    const std::vector<std::string>& operator[](size_t i) const { return p[i]; }
  } utts{all.data() + first, count};
  if (condition.stage > (size_t)jbk::kMglsaMaxStage)
    throw Error(kErrUnsupported, "stage (GAMMA option) > 8");
  if (condition.stage != 0 && voices_.stream_metadata(0).vector_length > (size_t)jbk::kMglsaMaxOrder)
    throw Error(kErrUnsupported, "stage != 0 with a spectrum vector longer than 36");
  if (condition.stage == 0 && (voices_.stream_metadata(0).vector_length < 2 || voices_.stream_metadata(0).vector_length > 128))
    throw Error(kErrUnsupported, "mel-cepstral vectors of 2..128 coefficients");
  if (voices_.global_metadata().num_streams < 3 || voices_.stream_metadata(2).vector_length == 0)
    throw Error(kErrUnsupported, "voices without an LPF stream are not built yet");
  std::vector<UtterancePack> packs(utts.size());
  const std::vector<StreamTable>* tabs = device_gather ? &tables_ : nullptr;
  size_t nthreads = std::min<size_t>(host_threads(), utts.size());
  if (nthreads <= 1) {
    for (size_t u = 0; u < utts.size(); ++u) packs[u] = build_utterance(voices_, condition, utts[u], tabs);
    return packs;
  }
  std::atomic<size_t> next{0};
  std::mutex emu;
  std::unique_ptr<Error> err;
  std::vector<std::thread> th;
  for (size_t t = 0; t < nthreads; ++t)
    th.emplace_back([&]() {
      for (;;) {
        size_t u = next.fetch_add(1);
        if (u >= utts.size()) return;
        try {
          packs[u] = build_utterance(voices_, condition, utts[u], tabs);
        } catch (const Error& e) {
          std::lock_guard<std::mutex> g(emu);
          if (!err) err.reset(new Error(e));
        } catch (const std::exception& e) {
          std::lock_guard<std::mutex> g(emu);
          if (!err) err.reset(new Error(kErrArgument, e.what()));
        }
      }
    });
  for (auto& t : th) t.join();
  if (err) throw *err;
  return packs;
}

// ---------------------------------------------------------------- resident batch
struct Engine::Resident {
  DeviceCtx* ctx = nullptr;
  int n_utt = 0;
  int64_t total_states = 0, total_frames = 0, total_samples = 0, voiced_samples = 0;
  int64_t max_frames = 0;
  std::vector<int64_t> frame_off;  // host copy, [n_utt+1]
  void* d_in = nullptr;
  size_t in_cap = 0, in_bytes = 0;
  void* d_work = nullptr;
  size_t work_cap = 0;
  jbk::BatchLayout layout{};
  jbk::StreamSet streams{};
  jbk::VocoderParams vp{};
  jbk::VocoderDev vd{};
  bool streaming = false;
  double* d_chunk = nullptr;  // streaming gather buffer
  double* mg_state = nullptr; // streaming with stage != 0: MGLSA delay lines [n_utt][8][36] carried between chunks
  size_t chunk_cap = 0;
  int64_t h2d_bytes = 0;
  // device-side gather of the state-level parameters (streams whose pdf table is resident)
  jbk::GatherStream gather[3]{};
  bool gathered[3] = {false, false, false};
  // MLSA variant 3: time segments (empty when the batch is streamed)
  jbk::MlsaSegments segs{};
  bool rows_given = false;         // prepare_tracks with coefficient_rows: vd.mcp holds the rows to copy into vd.b
  bool from_tracks = false;        // built by prepare_tracks: no MLPG inputs, run() starts at the vocoder
  jbk::MlsaSegments whole{};       // stage != 0: one segment per utterance (no cuts), for the repeat pass
  double lsp_beta = 0.0;           // stage != 0: beta of postfilter_lsp
  double* d_worst = nullptr;
  int64_t seg_frames = 0;          // frames filtered including warm-up
  std::vector<int32_t> wave_seg_first;  // [waves+1] segment range of every filter launch
  std::vector<int32_t> wave_utt_first;  // [waves+1] utterance range whose output that launch completes
  double beta = 0.0;               // > 0: post-filter kernels run after frame_prep
  bool used_fallback = false;      // the last pass re-ran the sequential kernel
  int repaired_segments = 0;       // the last pass re-filtered this many segments with a longer warm-up
  std::vector<jbk::MlsaSegment> seg_table;  // host copy of segs.seg (for the repair pass)
  int warm_frames = 0;             // warm-up the segments were cut with
  int warm_loud = 0;               // > 0: cuts behind loud frames were given this warm-up instead (mlsa_warm_loud)
  bool copied_by_waves = false;    // the last pass already queued its device-to-host copies
  cudaEvent_t kernels_end = nullptr;  // recorded after the last kernel of the last pass (variant 3)
  double* h_worst = nullptr;          // pinned word the last pass copies its cut-check result to
  int* h_fail = nullptr;              // pinned word: cuts of the last pass that failed the check
  double* out_direct = nullptr;       // output mode 2: the pinned host buffer the last pass stored into
  double cut_tol = 0.0;               // waveform-error bound the last pass checked its cuts against
  const double* result = nullptr;  // device waveform of the last pass
  // kernels per pass: expand, 3 solves (+ GV kernels), 2 frame preps, pulse (memset excluded),
  // excitation, MLSA (+ cut check for variant 3)
  int launches() const {
    if (from_tracks)
      return 2 + (rows_given ? 0 : 1) + (vd.binc ? 1 : 0) + 2 + (segs.n_seg > 0 ? (int)wave_seg_first.size() : 1) +
             (beta > 0.0 ? 2 : 0);
    int n = 1 + 3 + 2 + 4 + (segs.n_seg > 0 ? (int)wave_seg_first.size() - 1 : 0) + (beta > 0.0 ? 2 : 0) +
            (vd.binc ? 1 : 0);
    for (int i = 0; i < 3; ++i) n += gathered[i] ? 1 : 0;
    for (int i = 0; i < streams.n; ++i) n += streams.s[i].has_gv ? 1 : 0;
    // the 1-3-3 window shape runs as k_state_ivar + k_band_build + k_ldl instead of one solve kernel
    for (int i = 0; i < streams.n; ++i)
      n += (streams.s[i].nwin == 3 && streams.s[i].windows.width[1] == 3 && streams.s[i].windows.width[2] == 3) ? 2 : 0;
    return n;
  }
};

static jbk::WindowSet make_windows(const Windows& w) {
  jbk::WindowSet ws{};
  if (w.size() > (size_t)jbk::kMaxWin) throw Error(kErrUnsupported, "more than 4 windows per stream");
  ws.nwin = (int)w.size();
  for (size_t i = 0; i < w.size(); ++i) {
    if (w.windows[i].width() > (size_t)jbk::kMaxWinWidth)
      throw Error(kErrUnsupported, "window wider than 5 taps");
    ws.width[i] = (int)w.windows[i].width();
    for (size_t k = 0; k < w.windows[i].width(); ++k) ws.coef[i][k] = w.windows[i].coefficients[k];
  }
  return ws;
}

// Time segments of the filter kernels (MLSA variant 3, MGLSA): what make_resident uploads.
struct SegmentPlan {
  std::vector<jbk::MlsaSegment> segs;
  std::vector<int32_t> wave_seg_first, wave_utt_first;  // [waves + 1]
  int64_t seg_frames = 0;                               // frames filtered including warm-up
};

// Cuts every utterance into segments with W warm-up frames; `target` = how many segments fill the chip
// once (warps per scheduler x segments per warp x schedulers).
// `warm_at(u, begin)` (optional): warm-up of a cut at frame `begin` of utterance u when it is not W for all
// (Engine::mlsa_warm_loud); the cuts of an utterance are then placed so that all its segments still filter the same
// number of frames.
using WarmAt = std::function<int64_t(size_t, int64_t)>;
static SegmentPlan cut_segments(const std::vector<int64_t>& frame_off, size_t n_utt, int64_t W, int64_t target,
                                const std::vector<double>& wave_weights, const WarmAt& warm_at = WarmAt()) {
  SegmentPlan pl;
  std::vector<jbk::MlsaSegment>& segs = pl.segs;
  pl.wave_seg_first.assign(1, 0);
  pl.wave_utt_first.assign(1, 0);
  const int64_t TF = frame_off[n_utt];
  if (TF > 0) {
    std::vector<double> ww = wave_weights;
    if (ww.empty()) ww.assign(1, 1.0);
    // every wave should still fill the chip with segments of at least kSegMinFrames / 2 own frames:
    // merge the last two waves until the smallest one does
    auto smallest = [&]() {
      double s = 0, m = 1e300;
      for (double x : ww) s += x;
      for (double x : ww) m = std::min(m, x / s);
      return m;
    };
    while (ww.size() > 1 && smallest() * (double)TF < (double)(target * (kSegMinFrames / 2))) {
      ww[ww.size() - 2] += ww.back();
      ww.pop_back();
    }
    const int waves = (int)ww.size();
    double wsum = 0, wacc = 0;
    for (double x : ww) wsum += x;
    std::vector<size_t> ucut(waves + 1, n_utt);
    ucut[0] = 0;
    for (int wv = 1; wv < waves; ++wv) {
      wacc += ww[(size_t)wv - 1];
      size_t u = ucut[wv - 1];
      while (u < n_utt && (double)frame_off[u] < (double)TF * (wacc / wsum)) ++u;
      ucut[wv] = u;
    }
    for (int wv = 0; wv < waves; ++wv) {
      const size_t ua = ucut[wv], ub = ucut[wv + 1];
      const int64_t TFw = frame_off[ub] - frame_off[ua];
      if (TFw <= 0) continue;
      std::vector<jbk::MlsaSegment> ws;
      bool sized = false;  // the search for S below only counts segments: the warm-up shares are placed once, at the end
      auto cut = [&](int64_t S) {
        ws.clear();
        for (size_t u = ua; u < ub; ++u) {
          const int64_t T = frame_off[u + 1] - frame_off[u];
          if (T <= 0) continue;
          int64_t G = std::max<int64_t>(1, (T + S / 2) / S);
          // equal filtered length L for every segment: the first owns L frames, the others L - W
          while (G > 1 && (T + (G - 1) * W) / G - W < kSegMinFrames / 2) --G;
          if (warm_at && sized && G > 1) {
            // the share of a cut depends on where it falls and where it falls on the shares before it: iterate
            std::vector<int64_t> beg((size_t)G + 1, 0), wk((size_t)G, 0);
            int64_t sum_w = (G - 1) * W;
            bool ok = false;
            for (int it = 0; it < 6 && !ok; ++it) {
              const int64_t L = (T + sum_w + G - 1) / G;
              int64_t pos = std::min(L, T), sw = 0;
              for (int64_t k = 1; k < G; ++k) {
                const int64_t w = std::min(warm_at(u, pos), pos);
                beg[(size_t)k] = pos;
                wk[(size_t)k] = w;
                sw += w;
                pos = std::min(T, pos + std::max<int64_t>(L - w, kSegMinFrames / 2));
              }
              beg[(size_t)G] = T;
              ok = sw == sum_w;
              sum_w = sw;
            }
            bool fits = true;  // converged or not, the cuts are valid as long as every segment owns enough frames
            for (int64_t k = 0; k < G; ++k) fits = fits && beg[(size_t)k + 1] - beg[(size_t)k] >= kSegMinFrames / 2;
            if (fits) {
              for (int64_t k = 0; k < G; ++k) {
                jbk::MlsaSegment sg;
                sg.utt = (int32_t)u;
                sg.begin = (int32_t)beg[(size_t)k];
                sg.end = (int32_t)beg[(size_t)k + 1];
                sg.warm = (int32_t)(beg[(size_t)k] - wk[(size_t)k]);
                sg.prev = k == 0 ? -1 : (int32_t)(ws.size() - 1);
                ws.push_back(sg);
              }
              continue;
            }
          }
          const int64_t L = (T + (G - 1) * W + G - 1) / G;
          int64_t begin = 0;
          for (int64_t k = 0; k < G && begin < T; ++k) {
            int64_t end = (k == G - 1) ? T : std::min<int64_t>(T, begin + (k == 0 ? L : L - W));
            jbk::MlsaSegment sg;
            sg.utt = (int32_t)u;
            sg.begin = (int32_t)begin;
            sg.end = (int32_t)end;
            sg.warm = (int32_t)std::max<int64_t>(0, begin - W);
            sg.prev = k == 0 ? -1 : (int32_t)(ws.size() - 1);  // index before sorting, remapped below
            ws.push_back(sg);
            begin = end;
          }
        }
      };
      int64_t S = std::max<int64_t>(waves > 1 ? kSegMinFrames / 2 : kSegMinFrames, (TFw + target - 1) / target);
      cut(S);
      // One wave: all segments run about equally long, so a grid slightly larger than the chip holds
      // (target segments) would leave a second, nearly empty wave as long as the first.
      for (int it = 0; it < 64 && (int64_t)ws.size() > target && (int64_t)ws.size() < 2 * target; ++it) {
        S += std::max<int64_t>(1, S / 50);
        cut(S);
      }
      if (warm_at) {
        sized = true;
        cut(S);
      }
      // six consecutive segments share a warp and run as long as the longest of them
      std::vector<int32_t> perm(ws.size()), where(ws.size());
      for (size_t i = 0; i < ws.size(); ++i) perm[i] = (int32_t)i;
      std::stable_sort(perm.begin(), perm.end(), [&](int32_t x, int32_t y) {
        return ws[x].end - ws[x].warm > ws[y].end - ws[y].warm;
      });
      // Two blocks (of 4 warps x 6 segments) are resident per SM and block b shares its SM with block
      // b + #SMs: the second round of blocks takes the segments in ascending order, so that the longest
      // warps sit beside the shortest ones and every scheduler gets about the same number of frames.
      {
        const size_t first_round = (size_t)(target / std::max(1, kSegWarpsPerSmsp));  // #SMs x 24 segments
        if (kSegWarpsPerSmsp == 2 && perm.size() > first_round) std::reverse(perm.begin() + first_round, perm.end());
      }
      for (size_t i = 0; i < perm.size(); ++i) where[perm[i]] = (int32_t)i;
      const int32_t base = (int32_t)segs.size();
      if (wv > 0) {
        pl.wave_seg_first.push_back(base);
        pl.wave_utt_first.push_back((int32_t)ua);
      }
      for (size_t i = 0; i < perm.size(); ++i) {
        jbk::MlsaSegment sg = ws[perm[i]];
        if (sg.prev >= 0) sg.prev = base + where[sg.prev];
        pl.seg_frames += sg.end - sg.warm;
        segs.push_back(sg);
      }
    }
  }
  pl.wave_seg_first.push_back((int32_t)segs.size());
  pl.wave_utt_first.push_back((int32_t)n_utt);
  return pl;
}

std::vector<int32_t> Engine::plan_segments_debug(const std::vector<int64_t>& frames, int warm, int warm_loud,
                                                 double threshold, const double* gain, int64_t target) {
  const size_t n = frames.size();
  std::vector<int64_t> frame_off(n + 1, 0);
  for (size_t u = 0; u < n; ++u) frame_off[u + 1] = frame_off[u] + frames[u];
  std::vector<std::vector<double>> cum(n);
  WarmAt warm_at;
  if (gain && warm_loud > 0 && warm_loud < warm) {
    for (size_t u = 0; u < n; ++u) {
      cum[u].assign((size_t)frames[u] + 1, 0.0);
      for (int64_t f = 0; f < frames[u]; ++f) cum[u][(size_t)f + 1] = cum[u][(size_t)f] + gain[frame_off[u] + f];
    }
    const int64_t wl = warm_loud, wq = warm;
    const double thr = threshold * (double)wl;
    warm_at = [&cum, wl, wq, thr](size_t u, int64_t begin) -> int64_t {  // the rule of make_resident
      if (begin <= wl) return begin;
      return cum[u][(size_t)begin] - cum[u][(size_t)(begin - wl)] > thr ? wl : wq;
    };
  }
  const SegmentPlan pl = cut_segments(frame_off, n, warm, target, {}, warm_at);
  std::vector<int32_t> out;
  for (const jbk::MlsaSegment& s : pl.segs) {
    out.push_back(s.utt); out.push_back(s.begin); out.push_back(s.end); out.push_back(s.warm); out.push_back(s.prev);
  }
  return out;
}

// Batching layer: pack the state-level parameters of `packs` into one pinned blob, copy it to
// the device, and lay out every work array of the pipeline in one arena.
const std::vector<double>& Engine::row_gain(double alpha) const {
  std::lock_guard<std::mutex> g(row_gain_mu_);
  if (row_gain_alpha_ != alpha || row_gain_.empty()) {
    const StreamTable& t = tables_[0];
    const size_t D = voices_.stream_metadata(0).vector_length;
    row_gain_.assign(t.rows, 0.0);
    for (size_t r = 0; r < t.rows; ++r) {
      const double* c = t.mean.data() + r * t.width;
      double b = c[D - 1];  // mc2b (cepstrum.rs:139-149), coefficient 0 only
      for (size_t d = D - 1; d-- > 0;) b = c[d] - alpha * b;
      row_gain_[r] = b;
    }
    row_gain_alpha_ = alpha;
  }
  return row_gain_;
}

Engine::Resident* Engine::make_resident(const Engine& e, DeviceCtx* ctx, const UtterancePack* packs,
                                        size_t n_utt, bool streaming, const Condition& cond,
                                        const std::vector<double>& wave_weights) {
  std::unique_ptr<Engine::Resident> r(new Engine::Resident());
  r->ctx = ctx;
  r->n_utt = (int)n_utt;
  r->streaming = streaming;
  const VoiceSet& vs = e.voices();
  const size_t nstream = 3;
  const int fperiod = (int)cond.fperiod;

  std::vector<int64_t> state_off(n_utt + 1, 0);
  r->frame_off.assign(n_utt + 1, 0);
  for (size_t u = 0; u < n_utt; ++u) {
    state_off[u + 1] = state_off[u] + (int64_t)packs[u].nstates;
    r->frame_off[u + 1] = r->frame_off[u] + (int64_t)packs[u].nframes;
    r->max_frames = std::max<int64_t>(r->max_frames, (int64_t)packs[u].nframes);
  }
  const int64_t TS = r->total_states = state_off[n_utt];
  const int64_t TF = r->total_frames = r->frame_off[n_utt];
  r->total_samples = TF * fperiod;
  if (r->max_frames * (int64_t)fperiod > (int64_t)INT32_MAX)
    throw Error(kErrUnsupported, "utterance longer than 2^31 samples");

  // ---- input blob layout
  Carver in;
  const size_t o_soff = in.take((n_utt + 1) * 8), o_foff = in.take((n_utt + 1) * 8);
  const size_t o_dur = in.take((size_t)TS * 4);
  const size_t o_order = in.take(n_utt * 4);
  // MLSA variant 3: cut every utterance into time segments (kernels.h). The target is enough
  // segments for kSegWarpsPerSmsp warps of six on every warp scheduler; a segment is never shorter
  // than kSegMinFrames of its own, so small batches simply run fewer warps.
  // With more than one entry in `wave_weights` (Engine::synthesize_batch) the utterances form that many
  // groups, with frame counts in proportion to the weights; each group is cut into its own full wave of
  // segments and filtered by its own launch, so that the device-to-host copy of one group overlaps
  // the filter of the next.
  SegmentPlan plan;
  plan.wave_seg_first.assign(2, 0);
  plan.wave_utt_first = {0, (int32_t)n_utt};
  const int seg_slots = cond.stage != 0 ? 32 / (int)cond.stage : 6;  // segments per warp of the filter kernel
  // Warm-up by loudness (engine.h: mlsa_warm_loud). The host has no spectra yet, but it has what they are made of:
  // the static mean of every state's spectrum pdf. Its log gain (coefficient 0 after mc2b) per frame, as prefix
  // sums per utterance, tells speech from pauses well enough to size the warm-up of a cut; the device check and the
  // repair pass stay in charge of correctness.
  std::vector<std::vector<double>> loud_cum;
  r->warm_loud = 0;
  if (!streaming && TF > 0 && cond.stage == 0 && e.mlsa_warm_loud > 0 && e.mlsa_warm_loud < e.mlsa_warm_frames &&
      ctx->loud_strikes < 2) {
    const StreamModelMetadata& md0 = vs.stream_metadata(0);
    const size_t D = md0.vector_length, L0 = D * vs.stream_windows(0).size();
    const bool by_rows = e.device_gather && !e.tables_.empty() && e.tables_[0].regular;
    const std::vector<double>& wgt = cond.interporation_weight.parameter[0];
    const size_t nv = std::min(vs.size(), wgt.size());
    loud_cum.resize(n_utt);
    bool usable = D >= 2;
    for (size_t u = 0; u < n_utt && usable; ++u) {
      const UtterancePack& p = packs[u];
      if (p.streams.empty()) usable = false;
      else if (by_rows ? p.streams[0].rows.size() != p.nstates * nv : p.streams[0].mean.size() != p.nstates * L0) usable = false;
    }
    if (usable) {
      // mc2b is linear, so the gain of an interpolated mean is the interpolation of the rows' gains
      const std::vector<double>* rg = by_rows ? &e.row_gain(cond.alpha) : nullptr;
      parallel_for(n_utt, [&](size_t u) {
        const UtterancePack& p = packs[u];
        std::vector<double>& cum = loud_cum[u];
        cum.assign(p.nframes + 1, 0.0);
        size_t f = 0;
        for (size_t i = 0; i < p.nstates; ++i) {
          double b = 0.0;
          if (by_rows) {
            const int32_t* rr = p.streams[0].rows.data() + i * nv;
            for (size_t q = 0; q < nv; ++q) b += wgt[q] * (*rg)[(size_t)rr[q]];
          } else {
            const double* c = p.streams[0].mean.data() + i * L0;
            b = c[D - 1];  // mc2b (cepstrum.rs:139-149), coefficient 0 only
            for (size_t d = D - 1; d-- > 0;) b = c[d] - cond.alpha * b;
          }
          for (int32_t k = 0; k < p.durations[i] && f < p.nframes; ++k, ++f) cum[f + 1] = cum[f] + b;
        }
        for (; f < p.nframes; ++f) cum[f + 1] = cum[f];
      });
      r->warm_loud = e.mlsa_warm_loud;
    }
  }
  WarmAt warm_at;
  if (r->warm_loud > 0) {
    const int64_t wl = r->warm_loud, wq = e.mlsa_warm_frames;
    const double thr = e.mlsa_loud_threshold * (double)wl;
    warm_at = [&loud_cum, wl, wq, thr](size_t u, int64_t begin) -> int64_t {
      if (begin <= wl) return begin;  // reaches the start of the utterance, where the zero state is the true one
      const std::vector<double>& cum = loud_cum[u];
      return cum[(size_t)begin] - cum[(size_t)(begin - wl)] > thr ? wl : wq;
    };
  }
  if (!streaming && TF > 0)
    plan = cut_segments(r->frame_off, n_utt, e.mlsa_warm_frames * (cond.stage != 0 ? kMglsaWarmFactor : 1),
                        (int64_t)ctx->sm_count * 4 * kSegWarpsPerSmsp * seg_slots, wave_weights, warm_at);
  const std::vector<jbk::MlsaSegment>& segs = plan.segs;
  r->wave_seg_first = plan.wave_seg_first;
  r->wave_utt_first = plan.wave_utt_first;
  r->seg_frames = plan.seg_frames;
  const size_t o_segs = in.take(segs.size() * sizeof(jbk::MlsaSegment));
  // stage != 0: one segment per utterance, longest first, for the pass that repeats a failed cut check
  std::vector<jbk::MlsaSegment> whole;
  if (cond.stage != 0 && !streaming) {
    for (size_t u = 0; u < n_utt; ++u)
      if (packs[u].nframes > 0) whole.push_back(jbk::MlsaSegment{(int32_t)u, 0, (int32_t)packs[u].nframes, 0, -1});
    std::stable_sort(whole.begin(), whole.end(),
                     [](const jbk::MlsaSegment& x, const jbk::MlsaSegment& y) { return x.end > y.end; });
  }
  const size_t o_whole = in.take(whole.size() * sizeof(jbk::MlsaSegment));
  size_t o_mean[3], o_vari[3], o_msd[3], o_gvm[3], o_gvv[3], o_gsw[3], o_rows[3];
  int dim[3], nwin[3], is_msd[3], has_gv[3], nvoice[3];
  bool gath[3];
  for (size_t s = 0; s < nstream; ++s) {
    // device gather: row indices travel instead of the interpolated vectors
    gath[s] = e.device_gather && s < e.tables_.size() && e.tables_[s].regular;
    nvoice[s] = (int)std::min(vs.size(), cond.interporation_weight.parameter[s].size());
    const StreamModelMetadata& md = vs.stream_metadata(s);
    dim[s] = (int)md.vector_length;
    nwin[s] = (int)vs.stream_windows(s).size();
    is_msd[s] = md.is_msd;
    has_gv[s] = md.use_gv;
    const size_t L = (size_t)dim[s] * nwin[s];
    o_rows[s] = gath[s] ? in.take((size_t)TS * nvoice[s] * 4) : 0;
    o_mean[s] = gath[s] ? 0 : in.take((size_t)TS * L * 8);
    o_vari[s] = gath[s] ? 0 : in.take((size_t)TS * L * 8);
    o_msd[s] = (is_msd[s] && !gath[s]) ? in.take((size_t)TS * 8) : 0;
    o_gvm[s] = has_gv[s] ? in.take(n_utt * dim[s] * 8) : 0;
    o_gvv[s] = has_gv[s] ? in.take(n_utt * dim[s] * 8) : 0;
    o_gsw[s] = has_gv[s] ? in.take((size_t)TS) : 0;
  }
  r->in_bytes = in.off;
  uint8_t* h = (uint8_t*)pinned().alloc(r->in_bytes);
  struct Release {
    void* p;
    ~Release() { pinned().release(p); }
  } rel{h};
  memcpy(h + o_soff, state_off.data(), (n_utt + 1) * 8);
  memcpy(h + o_foff, r->frame_off.data(), (n_utt + 1) * 8);
  if (!segs.empty()) memcpy(h + o_segs, segs.data(), segs.size() * sizeof(jbk::MlsaSegment));
  if (!whole.empty()) memcpy(h + o_whole, whole.data(), whole.size() * sizeof(jbk::MlsaSegment));
  {  // longest utterance first: the per-utterance kernels are scheduled block by block in this order
    std::vector<int32_t> order(n_utt);
    for (size_t u = 0; u < n_utt; ++u) order[u] = (int32_t)u;
    std::stable_sort(order.begin(), order.end(),
                     [&](int32_t x, int32_t y) { return packs[x].nframes > packs[y].nframes; });
    if (n_utt) memcpy(h + o_order, order.data(), n_utt * 4);
  }
  std::atomic<int64_t> voiced_frames{0};
  auto pack_one = [&](size_t u) {
    const UtterancePack& p = packs[u];
    const size_t so = (size_t)state_off[u];
    if (p.nstates) memcpy(h + o_dur + so * 4, p.durations.data(), p.nstates * 4);
    int64_t vf = 0;
    for (size_t s = 0; s < nstream && s < p.streams.size(); ++s) {
      const StreamPack& sp = p.streams[s];
      const size_t L = (size_t)dim[s] * nwin[s];
      if (gath[s]) {
        if (sp.rows.size() != p.nstates * (size_t)nvoice[s])
          throw Error(kErrArgument, "stream pack has the wrong shape (row indices)");
        if (p.nstates) memcpy(h + o_rows[s] + so * nvoice[s] * 4, sp.rows.data(), sp.rows.size() * 4);
      } else if (sp.mean.size() != p.nstates * L) {
        throw Error(kErrArgument, "stream pack has the wrong shape");
      } else if (p.nstates) {
        memcpy(h + o_mean[s] + so * L * 8, sp.mean.data(), p.nstates * L * 8);
        memcpy(h + o_vari[s] + so * L * 8, sp.vari.data(), p.nstates * L * 8);
        if (is_msd[s]) memcpy(h + o_msd[s] + so * 8, sp.msd.data(), p.nstates * 8);
      }
      if (has_gv[s]) {
        double* gm = (double*)(h + o_gvm[s]) + u * dim[s];
        double* gv = (double*)(h + o_gvv[s]) + u * dim[s];
        for (int d = 0; d < dim[s]; ++d) {
          gm[d] = sp.has_gv ? sp.gv_mean[d] : 0.0;
          gv[d] = sp.has_gv ? sp.gv_vari[d] : 0.0;
        }
        if (p.nstates) {
          if (sp.has_gv) memcpy(h + o_gsw[s] + so, sp.gv_switch.data(), p.nstates);
          else memset(h + o_gsw[s] + so, 0, p.nstates);
        }
      }
      if (s == 1 && is_msd[s]) {  // voiced frames, for the flop count of the report only
        const std::vector<double>& w = cond.interporation_weight.parameter[s];
        for (size_t i = 0; i < p.nstates; ++i) {
          double m;
          if (gath[s]) {  // same arithmetic as k_gather_states / VoiceSet::weighted
            const int32_t* rr = sp.rows.data() + i * nvoice[s];
            m = w[0] * e.tables_[s].msd[(size_t)rr[0]];
            for (int q = 1; q < nvoice[s]; ++q) m += w[q] * e.tables_[s].msd[(size_t)rr[q]];
          } else {
            m = sp.msd[i];
          }
          if (m > cond.msd_threshold[1]) vf += p.durations[i];
        }
      }
    }
    voiced_frames += vf;
  };
  parallel_for(n_utt, pack_one);  // the batching layer copies ~0.5 MB per utterance: use every core
  r->voiced_samples = voiced_frames.load() * fperiod;

  JB_CUDA(cudaSetDevice(ctx->device));
  if (e.device_gather) ctx->load_tables(e.tables_);
  r->d_in = ctx->dalloc(r->in_bytes, &r->in_cap);
  JB_CUDA(cudaEventRecord(ctx->ev[0], ctx->stream));
  JB_CUDA(cudaMemcpyAsync(r->d_in, h, r->in_bytes, cudaMemcpyHostToDevice, ctx->stream));
  JB_CUDA(cudaEventRecord(ctx->ev[1], ctx->stream));
  r->h2d_bytes = (int64_t)r->in_bytes;

  // ---- work arena
  Carver wk;
  const size_t w_fstate = wk.take((size_t)TF * 4);
  size_t w_left[3] = {0}, w_right[3] = {0}, w_vidx[3] = {0}, w_nv[3] = {0}, w_par[3];
  size_t scratch_bytes = 0, scratch1_bytes = 0;  // streams 1/2 run one after the other on the side stream; stream 0 beside them
  for (size_t s = 0; s < nstream; ++s) {
    if (is_msd[s]) {
      w_left[s] = wk.take((size_t)TF * 4);
      w_right[s] = wk.take((size_t)TF * 4);
      w_vidx[s] = wk.take((size_t)TF * 4);
      w_nv[s] = wk.take(n_utt * 4);
    }
    w_par[s] = wk.take((size_t)TF * dim[s] * 8);
    jbk::WindowSet ws = make_windows(vs.stream_windows(s));
    size_t& sb = (s != 0) ? scratch1_bytes : scratch_bytes;
    sb = std::max(sb, (size_t)jbk::mlpg_scratch_planes(ws.band_width(), has_gv[s]) *
                                                (size_t)TF * dim[s] * 8);
  }
  size_t w_ivar[3] = {0};  // MeanVari::with_ivar per state, written by the MLPG split path
  for (size_t s = 0; s < nstream; ++s) w_ivar[s] = wk.take((size_t)TS * dim[s] * nwin[s] * 8);
  size_t w_gmean[3] = {0}, w_gvari[3] = {0}, w_gmsd[3] = {0};
  for (size_t s = 0; s < nstream; ++s)
    if (gath[s]) {
      const size_t L = (size_t)dim[s] * nwin[s];
      w_gmean[s] = wk.take((size_t)TS * L * 8);
      w_gvari[s] = wk.take((size_t)TS * L * 8);
      w_gmsd[s] = is_msd[s] ? wk.take((size_t)TS * 8) : 0;
    }
  const size_t w_scratch = wk.take(scratch_bytes);
  const size_t w_scratch1 = wk.take(scratch1_bytes);
  const size_t w_b = wk.take((size_t)TF * dim[0] * 8);
  const size_t w_pitch = wk.take((size_t)TF * 8);
  const size_t w_pulse = wk.take((size_t)r->total_samples * 8);
  const size_t w_runs = wk.take(((size_t)TF + 2) * 4);
  const size_t w_wave = wk.take((size_t)r->total_samples * 8);
  const size_t w_state = streaming ? wk.take(n_utt * jbk::kVocStateDoubles * 8) : 0;
  const size_t w_mgstate = (streaming && cond.stage != 0) ? wk.take(n_utt * (size_t)jbk::kMglsaMaxStage * jbk::kMglsaMaxOrder * 8) : 0;
  const size_t w_out = segs.empty() ? 0 : wk.take((size_t)r->total_samples * 8);
  const size_t seg_state_doubles =
      cond.stage != 0 ? (size_t)jbk::mglsa_state_doubles((int)cond.stage) : (size_t)jbk::kMlsaSegStateDoubles;
  const size_t w_segstate = segs.empty() ? 0 : wk.take(segs.size() * 2 * seg_state_doubles * 8);
  const size_t w_worst = wk.take(8);
  const size_t w_segmis = segs.empty() ? 0 : wk.take(segs.size() * 8);
  const size_t w_failcnt = wk.take(8);
  const size_t fail_cap = std::max<size_t>(64, segs.size() / 16);
  const size_t w_faillist = segs.empty() ? 0 : wk.take(fail_cap * 4);
  const bool postfilter = cond.stage == 0 && cond.beta > 0.0 && dim[0] > 2;  // cepstrum.rs:24
  const bool first_rows = postfilter || cond.stage != 0;  // rows the first frame starts from (mod.rs:79-109)
  const size_t w_bfirst = first_rows ? wk.take(n_utt * (size_t)dim[0] * 8) : 0;
  const bool need_binc = cond.stage != 0 || !segs.empty();  // the segmented filters read per-frame increments
  const size_t w_binc = need_binc ? wk.take((size_t)TF * dim[0] * 8) : 0;
  r->d_work = ctx->dalloc(wk.off, &r->work_cap);
  uint8_t* di = (uint8_t*)r->d_in;
  uint8_t* dw = (uint8_t*)r->d_work;

  r->layout.n_utt = (int)n_utt;
  r->layout.state_off = (const int64_t*)(di + o_soff);
  r->layout.frame_off = (const int64_t*)(di + o_foff);
  r->layout.durations = (const int32_t*)(di + o_dur);
  r->layout.frame_state = (int32_t*)(dw + w_fstate);
  r->layout.order = (const int32_t*)(di + o_order);
  r->layout.max_frames = (int)r->max_frames;
  r->streams.n = (int)nstream;
  for (size_t s = 0; s < nstream; ++s) {
    jbk::StreamDev& sd = r->streams.s[s];
    sd.dim = dim[s];
    sd.nwin = nwin[s];
    sd.is_msd = is_msd[s];
    sd.has_gv = has_gv[s];
    sd.windows = make_windows(vs.stream_windows(s));
    sd.msd_threshold = cond.msd_threshold[s];
    sd.gv_weight = cond.gv_weight[s];
    sd.mean = gath[s] ? (const double*)(dw + w_gmean[s]) : (const double*)(di + o_mean[s]);
    sd.vari = gath[s] ? (const double*)(dw + w_gvari[s]) : (const double*)(di + o_vari[s]);
    sd.ivar = (double*)(dw + w_ivar[s]);
    sd.msd = !is_msd[s] ? nullptr : gath[s] ? (const double*)(dw + w_gmsd[s]) : (const double*)(di + o_msd[s]);
    r->gathered[s] = gath[s];
    if (gath[s]) {
      jbk::GatherStream& gs = r->gather[s];
      gs.width = dim[s] * nwin[s];
      gs.is_msd = is_msd[s];
      gs.nvoices = nvoice[s];
      for (int q = 0; q < nvoice[s]; ++q) gs.w[q] = cond.interporation_weight.parameter[s][(size_t)q];
      gs.tab_mean = ctx->d_tab_mean[s];
      gs.tab_vari = ctx->d_tab_vari[s];
      gs.tab_msd = ctx->d_tab_msd[s];
      gs.rows = (const int32_t*)(di + o_rows[s]);
      gs.mean = (double*)(dw + w_gmean[s]);
      gs.vari = (double*)(dw + w_gvari[s]);
      gs.msd = is_msd[s] ? (double*)(dw + w_gmsd[s]) : nullptr;
      gs.half_tone = (s == 1 && cond.additional_half_tone != 0.0) ? 1 : 0;
      gs.half_tone_add = cond.additional_half_tone * kHalfTone;
      gs.lf0_min = kMinLf0;
      gs.lf0_max = kMaxLf0;
    }
    sd.gv_mean = has_gv[s] ? (const double*)(di + o_gvm[s]) : nullptr;
    sd.gv_vari = has_gv[s] ? (const double*)(di + o_gvv[s]) : nullptr;
    sd.gv_switch = has_gv[s] ? (const uint8_t*)(di + o_gsw[s]) : nullptr;
    sd.left = is_msd[s] ? (int32_t*)(dw + w_left[s]) : nullptr;
    sd.right = is_msd[s] ? (int32_t*)(dw + w_right[s]) : nullptr;
    sd.vidx = is_msd[s] ? (int32_t*)(dw + w_vidx[s]) : nullptr;
    sd.nvoiced = is_msd[s] ? (int32_t*)(dw + w_nv[s]) : nullptr;
    sd.scratch = (double*)(dw + (s != 0 ? w_scratch1 : w_scratch));
    sd.par = (double*)(dw + w_par[s]);
  }
  r->vp.nmcp = dim[0];
  r->vp.nlpf = dim[2];
  r->vp.fperiod = fperiod;
  r->vp.rate = (double)cond.sampling_frequency;
  r->vp.alpha = cond.alpha;
  r->vp.volume = cond.volume;
  r->vp.stage = (int)cond.stage;
  r->vp.use_log_gain = cond.use_log_gain ? 1 : 0;
  r->lsp_beta = cond.stage != 0 ? cond.beta : 0.0;
  r->vd.mcp = r->streams.s[0].par;
  r->vd.lf0 = r->streams.s[1].par;
  r->vd.lpf = r->streams.s[2].par;
  r->vd.b = (double*)(dw + w_b);
  r->vd.b_first = first_rows ? (const double*)(dw + w_bfirst) : nullptr;
  r->vd.binc = need_binc ? (double*)(dw + w_binc) : nullptr;
  r->beta = postfilter ? cond.beta : 0.0;
  r->vd.pitch = (double*)(dw + w_pitch);
  r->vd.pulse = (double*)(dw + w_pulse);
  r->vd.runs = (int32_t*)(dw + w_runs);
  r->vd.wave = (double*)(dw + w_wave);
  r->vd.state = streaming ? (double*)(dw + w_state) : nullptr;
  r->mg_state = (streaming && cond.stage != 0) ? (double*)(dw + w_mgstate) : nullptr;
  r->vd.out = segs.empty() ? nullptr : (double*)(dw + w_out);
  r->segs.n_seg = (int)segs.size();
  r->segs.seg = (const jbk::MlsaSegment*)(di + o_segs);
  r->segs.state = segs.empty() ? nullptr : (double*)(dw + w_segstate);
  r->segs.mismatch = segs.empty() ? nullptr : (double*)(dw + w_segmis);
  r->segs.fail_count = segs.empty() ? nullptr : (int*)(dw + w_failcnt);
  r->segs.fail_list = segs.empty() ? nullptr : (int*)(dw + w_faillist);
  r->segs.fail_cap = segs.empty() ? 0 : (int)fail_cap;
  r->seg_table = segs;
  r->warm_frames = (int)(e.mlsa_warm_frames * (cond.stage != 0 ? kMglsaWarmFactor : 1));
  r->whole.n_seg = (int)whole.size();
  r->whole.seg = (const jbk::MlsaSegment*)(di + o_whole);
  r->whole.state = nullptr;
  r->d_worst = (double*)(dw + w_worst);
  r->result = r->vd.wave;
  r->vd.frame_begin = 0;
  r->vd.frame_count = -1;

  // ---- global noise table on this device (F7: one sequence shared by every utterance)
  const size_t need = (size_t)r->max_frames * fperiod;
  if (need > ctx->noise_len) {
    std::lock_guard<std::mutex> g(e.noise_mu_);
    const std::vector<double>& tbl = e.noise_table(need + need / 2);
    if (ctx->d_noise) JB_CUDA(cudaFree(ctx->d_noise));
    JB_CUDA(cudaMalloc((void**)&ctx->d_noise, tbl.size() * 8));
    JB_CUDA(cudaMemcpy(ctx->d_noise, tbl.data(), tbl.size() * 8, cudaMemcpyHostToDevice));
    ctx->noise_len = tbl.size();
  }
  r->vd.noise = ctx->d_noise;
  // the staging blob may be recycled only once the copy has finished
  JB_CUDA(cudaEventSynchronize(ctx->ev[1]));
  return r.release();
}

// Spectrum frames -> filter coefficient rows vd.b (and vd.b_first).
static void launch_spectrum_prep(Engine::Resident* r, cudaStream_t st) {
  if (r->vp.stage != 0) {  // Stage::NonZero: LSP frames (vocoder/mod.rs:92-107,148-155)
    JB_LAUNCH(jbk::launch_lsp_prep(r->layout, r->vp, r->vd, r->lsp_beta, const_cast<double*>(r->vd.b_first),
                                   r->total_frames, st));
    JB_LAUNCH(jbk::launch_row_increments(r->layout, r->vp, r->vd, r->total_frames, st));
    return;
  }
  if (r->beta <= 0.0 && r->vd.binc && r->vd.b_first == nullptr && jbk::coef_rows_supported(r->vp)) {
    JB_LAUNCH(jbk::launch_coef_rows(r->layout, r->vp, r->vd, r->total_frames, st));  // rows and increments fused
    return;
  }
  JB_LAUNCH(jbk::launch_frame_prep(r->layout, r->vp, r->vd, r->total_frames, 2, st));
  if (r->beta > 0.0)  // MelCepstrum::postfilter_mcp (cepstrum.rs:23-37): rewrites vd.b in place
    JB_LAUNCH(jbk::launch_postfilter(r->layout, r->vp, r->vd, r->beta, const_cast<double*>(r->vd.b_first),
                                     r->total_frames, st));
  if (r->vd.binc) JB_LAUNCH(jbk::launch_row_increments(r->layout, r->vp, r->vd, r->total_frames, st));
}

// MLPG for the three streams + the per-frame vocoder inputs. The LF0 chain (solve, GV, pitch
// period and, for whole-utterance synthesis, the pulse train) only has n_utt systems, so it runs
// on the side stream beside the MCP/LPF solves instead of in front of them.
static void launch_mlpg_all(Engine::Resident* r, bool with_pulse) {
  DeviceCtx* c = r->ctx;
  cudaStream_t st = c->stream, sb = c->stream2;
  for (int s = 0; s < r->streams.n; ++s)
    if (r->gathered[s]) JB_LAUNCH(jbk::launch_gather_states(r->gather[s], r->total_states, st));
  JB_LAUNCH(jbk::launch_expand(r->layout, r->streams, st));
  JB_CUDA(cudaEventRecord(c->evx[0], st));
  JB_CUDA(cudaStreamWaitEvent(sb, c->evx[0], 0));
  JB_LAUNCH(jbk::launch_mlpg(r->layout, r->streams.s[1], r->total_frames, r->total_states, sb));
  JB_LAUNCH(jbk::launch_frame_prep(r->layout, r->vp, r->vd, r->total_frames, 1, sb));
  if (with_pulse) JB_LAUNCH(jbk::launch_pulse(r->layout, r->vp, r->vd, r->total_samples, sb));
  // the LPF stream (static window only: a per-frame kernel) needs nothing from MCP: it fills the side stream's slack
  JB_LAUNCH(jbk::launch_mlpg(r->layout, r->streams.s[2], r->total_frames, r->total_states, sb));
  JB_CUDA(cudaEventRecord(c->evx[1], sb));
  JB_LAUNCH(jbk::launch_mlpg(r->layout, r->streams.s[0], r->total_frames, r->total_states, st));
  launch_spectrum_prep(r, st);
  JB_CUDA(cudaStreamWaitEvent(st, c->evx[1], 0));
}

// Where the waveform goes while the filter is still running: wave w's utterance range is copied to
// host + its sample offset on `copy_stream` as soon as launch w has finished.
struct WaveCopy {
  int16_t* host16 = nullptr;  // 16-bit PCM delivery: quantise on the device (into dev16), copy 2 B per sample
  int16_t* dev16 = nullptr;
  double* host = nullptr;  // destination of utterance 0 of this resident batch
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t done = nullptr;  // recorded after the last copy (or after the filter when `direct`)
  bool direct = false;         // the filter stores straight into `host` (pinned, device-visible)
};

static void launch_vocoder(Engine::Resident* r, int variant, bool stop_after_excitation,
                           cudaEvent_t after_exc, bool pulse_done, double cut_tol, const WaveCopy* wc = nullptr) {
  cudaStream_t st = r->ctx->stream;
  // the noise table of the device may have been re-allocated (grown) by a batch prepared after this one:
  // take the pointer now, under the device mutex, not the one cached at prepare time
  r->vd.noise = r->ctx->d_noise;
  if (!pulse_done) JB_LAUNCH(jbk::launch_pulse(r->layout, r->vp, r->vd, r->total_samples, st));
  JB_LAUNCH(jbk::launch_excitation(r->layout, r->vp, r->vd, r->total_frames, st));
  if (after_exc) JB_CUDA(cudaEventRecord(after_exc, st));
  r->result = r->vd.wave;
  r->used_fallback = false;
  r->repaired_segments = 0;
  r->copied_by_waves = false;
  r->kernels_end = nullptr;
  r->out_direct = nullptr;
  r->cut_tol = cut_tol;
  if (stop_after_excitation) return;
  const bool mg = r->vp.stage != 0;  // MGLSA (mglsa.cu): segmented unless the sequential variant is forced
  if (mg && (r->vd.out == nullptr || r->vp.fperiod % 16 != 0 || r->streaming)) {
    // generate_step with stage != 0 (state carried in mg_state), or a frame period the segmented kernel does not
    // take: the sequential kernel, in place
    JB_LAUNCH(jbk::launch_mglsa_stream(r->layout, r->vp, r->vd, r->streaming ? r->mg_state : nullptr, st));
    return;
  }
  const bool v3 = mg ? (variant < 0 || variant == 3) && r->segs.n_seg > 0
                     : (variant < 0 || variant == 3) && r->segs.n_seg > 0 && jbk::mlsa_v3_supports(r->vp.nmcp) &&
                           jbk::mlsa_v3_chunk(r->vp.fperiod) > 0 && r->vd.frame_begin == 0 && r->vd.frame_count < 0;
  if (v3) {
    const int waves = (int)r->wave_seg_first.size() - 1;
    const bool direct = wc && wc->direct && !wc->host16;
    jbk::VocoderDev vd = r->vd;
    if (direct) {  // zero-copy: output stores cross PCIe while the filter runs
      vd.out = wc->host;
      r->out_direct = wc->host;
    }
    for (int w = 0; w < waves; ++w) {
      const int s0 = r->wave_seg_first[w], s1 = r->wave_seg_first[w + 1];
      if (mg) JB_LAUNCH(jbk::launch_mglsa(r->layout, r->vp, vd, r->segs, s0, s1 - s0, false, st));
      else JB_LAUNCH(jbk::launch_mlsa_v3(r->layout, r->vp, vd, r->segs, s0, s1 - s0, st));
      if (wc && !direct) {
        const int64_t f0 = r->frame_off[(size_t)r->wave_utt_first[w]];
        const int64_t f1 = r->frame_off[(size_t)r->wave_utt_first[w + 1]];
        const int64_t n0 = f0 * r->vp.fperiod, n1 = f1 * r->vp.fperiod;
        if (wc->host16) JB_LAUNCH(jbk::launch_pcm16(r->vd.out + n0, wc->dev16 + n0, n1 - n0, st));
        JB_CUDA(cudaEventRecord(r->ctx->evw[w & 1], st));
        JB_CUDA(cudaStreamWaitEvent(wc->copy_stream, r->ctx->evw[w & 1], 0));
        if (n1 > n0 && wc->host16)
          JB_CUDA(cudaMemcpyAsync(wc->host16 + n0, wc->dev16 + n0, (size_t)(n1 - n0) * 2,
                                  cudaMemcpyDeviceToHost, wc->copy_stream));
        else if (n1 > n0)
          JB_CUDA(cudaMemcpyAsync(wc->host + n0, r->vd.out + n0, (size_t)(n1 - n0) * 8,
                                  cudaMemcpyDeviceToHost, wc->copy_stream));
      }
    }
    r->h_worst = r->ctx->h_worst;
    r->h_fail = r->ctx->h_fail;
    r->segs.host_fail_count = r->h_fail;
    if (mg) {
      const int n = jbk::mglsa_state_doubles(r->vp.stage);
      JB_LAUNCH(jbk::launch_mlsa_seg_check(r->segs, r->d_worst, r->h_worst, st, n, n, true));
    } else {
      // the mismatch of every cut is scaled to a bound on the waveform error (kernels.h) and compared with cut_tol
      const jbk::SegCheckGain gain{&r->layout, r->vd.b, r->vp.nmcp, r->vp.volume, cut_tol};
      JB_LAUNCH(jbk::launch_mlsa_seg_check(r->segs, r->d_worst, r->h_worst, st, jbk::kMlsaSegStateDoubles, 5 * 37, false,
                                           nullptr, &gain));
    }
    JB_CUDA(cudaEventRecord(r->ctx->ev[7], st));
    r->kernels_end = r->ctx->ev[7];
    r->result = r->vd.out;
    if (wc) {
      JB_CUDA(cudaEventRecord(wc->done, direct ? st : wc->copy_stream));
      r->copied_by_waves = true;
    }
  } else if (mg) {
    // variants 0-2: whole utterances; 0 = the reference's operation order
    JB_LAUNCH(jbk::launch_mglsa(r->layout, r->vp, r->vd, r->whole, 0, r->whole.n_seg, variant == 0, st));
    r->result = r->vd.out;
  } else {
    JB_LAUNCH(jbk::launch_mlsa(r->layout, r->vp, r->vd, variant == 3 ? -1 : variant, st));
  }
}

// Variant 3 epilogue, after the stream has been synchronised: when the two sides of some cut
// disagree (the warm-up was too short for this input) the sequential kernel redoes the batch.
// Returns true when it did; the caller then repeats whatever it had queued behind the filter.
static bool mlsa_cut_check_failed(Engine::Resident* r, double tol) {
  if (r->result != r->vd.out || r->vd.out == nullptr || r->kernels_end == nullptr) return false;
  // MGLSA: the mismatch is relative to the state's scale (launch_mlsa_seg_check); held to 1e-9, an order
  // below what the LSP transform's conditioning leaves of the coefficient rows themselves
  if (r->vp.stage != 0) tol = std::min(tol, kMglsaCutRelTolerance);
  else if (r->h_fail && *r->h_fail > 0) return true;  // the device compared every cut's bound with cut_tol
  const double w = r->h_worst ? *r->h_worst : 0.0;
  return !(w <= tol);  // NaN counts as a failed cut
}
// A failed cut check, first resort (MLSA): the head of every segment behind a failing cut is filtered again from
// the exact state - the end state of the segment before it, which the original pass stored for the check - over
// `repair_frames` frames, into a scratch buffer (the pulse train's, consumed by now). Where the last re-filtered
// frame agrees with the original pass the two trajectories have merged: the new samples replace the old ones and
// the rest of the segment stands. The failing cuts come compacted from the device (MlsaSegments::fail_list).
// `ranges` receives (utterance, first sample, one past the last sample) of what changed, relative to the
// utterance. Returns false when that is not possible or not enough (many failing cuts, or a trajectory that has
// not merged): the caller then redoes the whole batch without cuts.
struct RepairRange {
  int32_t utt;
  int64_t n0, n1;
};
static bool mglsa_repair_segments(Engine::Resident* r, double tol, std::vector<RepairRange>* ranges);
static bool mlsa_repair_segments(Engine::Resident* r, double tol, std::vector<RepairRange>* ranges) {
  if (r->vp.stage != 0) return mglsa_repair_segments(r, tol, ranges);
  DeviceCtx* c = r->ctx;
  cudaStream_t st = c->stream;
  const int n = r->segs.n_seg;
  if (n == 0 || r->segs.fail_list == nullptr || (int)r->seg_table.size() != n || r->h_fail == nullptr) return false;
  const int nfail = *r->h_fail;
  if (r->warm_loud > 0 && nfail > std::max(4, n / 200) && c->loud_strikes < 2) ++c->loud_strikes;  // the rule does not fit this voice
  if (nfail <= 0 || nfail > r->segs.fail_cap || nfail > std::max(16, n / 16)) return false;
  std::vector<int32_t> ids((size_t)nfail);
  JB_CUDA(cudaMemcpyAsync(ids.data(), r->segs.fail_list, (size_t)nfail * 4, cudaMemcpyDeviceToHost, st));
  JB_CUDA(cudaStreamSynchronize(st));
  std::sort(ids.begin(), ids.end());
  // How long the mismatch left by the warm-up needs to fade below the merge tolerance: the slowest decay seen on
  // the shipped voices is ~0.2 decades per frame; a lone warp filters a frame in ~80 us, so the
  // range is kept as short as the worst bound of this pass allows. The merge check decides whether it was enough.
  const double worst = r->h_worst ? *r->h_worst : 0.0;
  int repair_frames = std::max(32, r->warm_frames);
  if (worst > 0.0 && worst < 1e300 && tol > 0.0) {
    const double decades = std::log10(std::max(worst / (0.2 * tol), 1.0));
    repair_frames = std::min(repair_frames, std::max(8, (int)std::ceil(decades / 0.2) + 4));
  }
  std::vector<jbk::MlsaSegment> fix;
  for (int32_t id : ids) {
    if (id < 0 || id >= n) return false;
    jbk::MlsaSegment s = r->seg_table[(size_t)id];  // `prev` keeps indexing the original pass's states
    if (s.prev < 0) return false;
    s.warm = s.begin;  // no warm-up: the kernel loads the end state of `prev`
    s.end = std::min<int32_t>(s.end, s.begin + repair_frames);
    fix.push_back(s);
  }
  std::stable_sort(fix.begin(), fix.end(), [](const jbk::MlsaSegment& a, const jbk::MlsaSegment& b) {
    return a.end - a.warm > b.end - b.warm;
  });
  Carver wk;
  const size_t o_tab = wk.take(fix.size() * sizeof(jbk::MlsaSegment));
  const size_t o_ok = wk.take(fix.size() * 4);
  size_t cap = 0;
  uint8_t* d = (uint8_t*)c->dalloc(wk.off, &cap);
  std::vector<int32_t> ok(fix.size(), 0);
  bool all_ok = false;
  try {
    JB_CUDA(cudaMemcpyAsync(d + o_tab, fix.data(), fix.size() * sizeof(jbk::MlsaSegment), cudaMemcpyHostToDevice, st));
    jbk::MlsaSegments sg{};
    sg.n_seg = (int)fix.size();
    sg.seg = (const jbk::MlsaSegment*)(d + o_tab);
    sg.init_state = r->segs.state;
    jbk::VocoderDev vd = r->vd;
    vd.out = r->vd.pulse;  // scratch, indexed like the output
    JB_LAUNCH(jbk::launch_mlsa_v3(r->layout, r->vp, vd, sg, 0, sg.n_seg, st));
    double* out = r->out_direct ? r->out_direct : r->vd.out;
    // merged when the last re-filtered frame agrees to a fifth of the waveform bound
    JB_LAUNCH(jbk::launch_mlsa_repair_merge(r->layout, sg.seg, sg.n_seg, vd.out, out, r->vp.fperiod, 0.2 * tol,
                                            (int*)(d + o_ok), st));
    JB_CUDA(cudaMemcpyAsync(ok.data(), d + o_ok, fix.size() * 4, cudaMemcpyDeviceToHost, st));
    JB_CUDA(cudaStreamSynchronize(st));  // `fix` and the scratch stay alive until here
    all_ok = true;
    for (int32_t x : ok) all_ok = all_ok && x == 1;
  } catch (...) {
    cudaStreamSynchronize(st);
    c->dfree(d, cap);
    throw;
  }
  c->dfree(d, cap);
  if (!all_ok) return false;
  r->repaired_segments = (int)fix.size();
  if (ranges) {
    ranges->clear();
    if (!r->out_direct)  // zero-copy output: the merge kernel has already stored into the host buffer
      for (const jbk::MlsaSegment& s : fix)
        ranges->push_back(RepairRange{s.utt, (int64_t)s.begin * r->vp.fperiod, (int64_t)s.end * r->vp.fperiod});
  }
  return true;
}

// Stage != 0 (MGLSA): the segments whose cut disagrees are filtered again with twice the warm-up and checked
// against the same end states; the rest of the batch stands.
static bool mglsa_repair_segments(Engine::Resident* r, double tol, std::vector<RepairRange>* ranges) {
  DeviceCtx* c = r->ctx;
  cudaStream_t st = c->stream;
  const int n = r->segs.n_seg;
  if (n == 0 || r->segs.mismatch == nullptr || (int)r->seg_table.size() != n || r->h_worst == nullptr) return false;
  if (r->out_direct) return false;
  tol = std::min(tol, kMglsaCutRelTolerance);
  std::vector<double> mis((size_t)n);
  JB_CUDA(cudaMemcpy(mis.data(), r->segs.mismatch, (size_t)n * 8, cudaMemcpyDeviceToHost));
  std::vector<jbk::MlsaSegment> fix;
  for (int i = 0; i < n; ++i)
    if (r->seg_table[(size_t)i].prev >= 0 && !(mis[(size_t)i] <= tol)) {
      jbk::MlsaSegment s = r->seg_table[(size_t)i];  // `prev` keeps indexing the original pass's states
      s.warm = std::max<int32_t>(0, s.begin - kRepairWarmFactor * std::max(1, r->warm_frames));
      fix.push_back(s);
    }
  if (fix.empty() || (int)fix.size() > std::max(16, n / 16)) return false;
  std::stable_sort(fix.begin(), fix.end(), [](const jbk::MlsaSegment& a, const jbk::MlsaSegment& b) {
    return a.end - a.warm > b.end - b.warm;
  });
  const size_t stride = (size_t)jbk::mglsa_state_doubles(r->vp.stage);
  Carver wk;
  const size_t o_tab = wk.take(fix.size() * sizeof(jbk::MlsaSegment));
  const size_t o_state = wk.take(fix.size() * 2 * stride * 8);
  size_t cap = 0;
  uint8_t* d = (uint8_t*)c->dalloc(wk.off, &cap);
  bool ok = false;
  try {
    JB_CUDA(cudaMemcpyAsync(d + o_tab, fix.data(), fix.size() * sizeof(jbk::MlsaSegment), cudaMemcpyHostToDevice, st));
    jbk::MlsaSegments sg{};
    sg.n_seg = (int)fix.size();
    sg.seg = (const jbk::MlsaSegment*)(d + o_tab);
    sg.state = (double*)(d + o_state);
    sg.mismatch = nullptr;
    JB_LAUNCH(jbk::launch_mglsa(r->layout, r->vp, r->vd, sg, 0, sg.n_seg, false, st));
    JB_LAUNCH(jbk::launch_mlsa_seg_check(sg, r->d_worst, r->h_worst, st, (int)stride, (int)stride, true, r->segs.state));
    JB_CUDA(cudaStreamSynchronize(st));  // `fix` and the scratch stay alive until here
    ok = *r->h_worst <= tol;
  } catch (...) {
    cudaStreamSynchronize(st);
    c->dfree(d, cap);
    throw;
  }
  c->dfree(d, cap);
  if (!ok) return false;
  r->repaired_segments = (int)fix.size();
  if (ranges) {
    ranges->clear();
    for (const jbk::MlsaSegment& s : fix)
      ranges->push_back(RepairRange{s.utt, (int64_t)s.begin * r->vp.fperiod, (int64_t)s.end * r->vp.fperiod});
  }
  return true;
}

static void mlsa_sequential_rerun(Engine::Resident* r) {
  cudaStream_t st = r->ctx->stream;
  if (r->vp.stage != 0) {
    JB_LAUNCH(jbk::launch_mglsa(r->layout, r->vp, r->vd, r->whole, 0, r->whole.n_seg, false, st));
    r->result = r->vd.out;
  } else {
    JB_LAUNCH(jbk::launch_mlsa(r->layout, r->vp, r->vd, 2, st));
    r->result = r->vd.wave;
  }
  r->used_fallback = true;
}

Engine::Resident* Engine::prepare(const std::vector<std::vector<std::string>>& utts,
                                  int device_index) const {
  std::vector<UtterancePack> packs = build_packs(utts);
  if (device_index < 0 || (size_t)device_index >= devices_.size())
    throw Error(kErrArgument, "device index out of range");
  DeviceCtx* c = ctx(devices_[(size_t)device_index]);
  std::lock_guard<std::mutex> g(c->mu);
  return make_resident(*this, c, packs.data(), packs.size(), false, condition, {});
}

Engine::Resident* Engine::prepare_tracks(const VocoderConfig& cfg, size_t n_utt, const size_t* frame_off,
                                         const double* spectrum, const double* lf0, const double* lpf,
                                         int device_index) const {
  if (device_index < 0 || (size_t)device_index >= devices_.size())
    throw Error(kErrArgument, "device index out of range");
  if (cfg.nmcp < 1 || cfg.fperiod < 1 || cfg.fperiod > 1024 || cfg.nlpf > 63)
    throw Error(kErrArgument, "vocoder: nmcp >= 1, 1 <= fperiod <= 1024, nlpf <= 63");
  if (cfg.nlpf > 0 && cfg.nlpf % 2 == 0)
    throw Error(kErrArgument, "The number of low-pass filter coefficient must be odd numbers.");  // speech.rs:38-40
  if (cfg.stage > (size_t)jbk::kMglsaMaxStage || (cfg.stage != 0 && (cfg.nmcp > (size_t)jbk::kMglsaMaxOrder || cfg.nmcp < 3)))
    throw Error(kErrUnsupported, "vocoder: stage <= 8, and 3 <= nmcp <= 36 when stage != 0");
  if (cfg.stage == 0 && (cfg.nmcp < 2 || cfg.nmcp > 128))
    throw Error(kErrUnsupported, "vocoder: mel-cepstral vectors of 2..128 coefficients");
  DeviceCtx* ctx = this->ctx(devices_[(size_t)device_index]);
  std::lock_guard<std::mutex> g(ctx->mu);
  JB_CUDA(cudaSetDevice(ctx->device));
  std::unique_ptr<Resident> r(new Resident());
  r->ctx = ctx;
  r->n_utt = (int)n_utt;
  r->from_tracks = true;
  r->frame_off.assign(n_utt + 1, 0);
  for (size_t u = 0; u < n_utt; ++u) {
    if (frame_off[u + 1] < frame_off[u] || frame_off[0] != 0) throw Error(kErrArgument, "vocoder: frame_off must ascend from 0");
    r->frame_off[u + 1] = (int64_t)frame_off[u + 1];
    r->max_frames = std::max<int64_t>(r->max_frames, (int64_t)(frame_off[u + 1] - frame_off[u]));
  }
  const int64_t TF = r->total_frames = r->frame_off[n_utt];
  const int fp = (int)cfg.fperiod;
  r->total_samples = TF * fp;
  if (r->max_frames * (int64_t)fp > (int64_t)INT32_MAX) throw Error(kErrUnsupported, "utterance longer than 2^31 samples");
  const size_t M = cfg.nmcp, NL = cfg.nlpf;

  const int seg_slots = cfg.stage != 0 ? 32 / (int)cfg.stage : 6;
  SegmentPlan plan;
  plan.wave_seg_first.assign(2, 0);
  plan.wave_utt_first = {0, (int32_t)n_utt};
  // Warm-up by loudness (make_resident): here the host holds the spectra themselves, so the log gain of a frame is
  // coefficient 0 of mc2b of its own vector. (Not with coefficient rows given: that is the test seam.)
  std::vector<std::vector<double>> loud_cum;
  WarmAt warm_at;
  r->warm_loud = 0;
  if (TF > 0 && cfg.stage == 0 && !cfg.coefficient_rows && M >= 2 && mlsa_warm_loud > 0 && mlsa_warm_loud < mlsa_warm_frames &&
      ctx->loud_strikes < 2) {
    loud_cum.resize(n_utt);
    const double alpha = cfg.alpha;
    parallel_for(n_utt, [&](size_t u) {
      const int64_t f0 = r->frame_off[u], T = r->frame_off[u + 1] - f0;
      std::vector<double>& cum = loud_cum[u];
      cum.assign((size_t)T + 1, 0.0);
      for (int64_t f = 0; f < T; ++f) {
        const double* c = spectrum + (size_t)(f0 + f) * M;
        double b = c[M - 1];  // mc2b (cepstrum.rs:139-149), coefficient 0 only
        for (size_t d = M - 1; d-- > 0;) b = c[d] - alpha * b;
        cum[(size_t)f + 1] = cum[(size_t)f] + b;
      }
    });
    r->warm_loud = mlsa_warm_loud;
    const int64_t wl = mlsa_warm_loud, wq = mlsa_warm_frames;
    const double thr = mlsa_loud_threshold * (double)wl;
    warm_at = [&loud_cum, wl, wq, thr](size_t u, int64_t begin) -> int64_t {
      if (begin <= wl) return begin;
      const std::vector<double>& cum = loud_cum[u];
      return cum[(size_t)begin] - cum[(size_t)(begin - wl)] > thr ? wl : wq;
    };
  }
  if (TF > 0)
    plan = cut_segments(r->frame_off, n_utt, mlsa_warm_frames * (cfg.stage != 0 ? kMglsaWarmFactor : 1),
                        (int64_t)ctx->sm_count * 4 * kSegWarpsPerSmsp * seg_slots, {}, warm_at);
  r->wave_seg_first = plan.wave_seg_first;
  r->wave_utt_first = plan.wave_utt_first;
  r->seg_frames = plan.seg_frames;
  std::vector<jbk::MlsaSegment> whole;
  std::vector<int32_t> order(n_utt);
  for (size_t u = 0; u < n_utt; ++u) order[u] = (int32_t)u;
  std::stable_sort(order.begin(), order.end(), [&](int32_t x, int32_t y) {
    return r->frame_off[(size_t)x + 1] - r->frame_off[(size_t)x] > r->frame_off[(size_t)y + 1] - r->frame_off[(size_t)y];
  });
  if (cfg.stage != 0)
    for (int32_t u : order) {
      const int64_t T = r->frame_off[(size_t)u + 1] - r->frame_off[(size_t)u];
      if (T > 0) whole.push_back(jbk::MlsaSegment{u, 0, (int32_t)T, 0, -1});
    }

  Carver in;
  const size_t o_foff = in.take((n_utt + 1) * 8), o_order = in.take(n_utt * 4);
  const size_t o_segs = in.take(plan.segs.size() * sizeof(jbk::MlsaSegment));
  const size_t o_whole = in.take(whole.size() * sizeof(jbk::MlsaSegment));
  const size_t o_spec = in.take(((size_t)TF + (cfg.coefficient_rows ? n_utt : 0)) * M * 8), o_lf0 = in.take((size_t)TF * 8), o_lpf = in.take((size_t)TF * NL * 8);
  r->in_bytes = in.off;
  uint8_t* h = (uint8_t*)pinned().alloc(r->in_bytes);
  struct Release {
    void* p;
    ~Release() { pinned().release(p); }
  } rel{h};
  memcpy(h + o_foff, r->frame_off.data(), (n_utt + 1) * 8);
  if (n_utt) memcpy(h + o_order, order.data(), n_utt * 4);
  if (!plan.segs.empty()) memcpy(h + o_segs, plan.segs.data(), plan.segs.size() * sizeof(jbk::MlsaSegment));
  if (!whole.empty()) memcpy(h + o_whole, whole.data(), whole.size() * sizeof(jbk::MlsaSegment));
  if (TF > 0) {
    memcpy(h + o_spec, spectrum, ((size_t)TF + (cfg.coefficient_rows ? n_utt : 0)) * M * 8);
    memcpy(h + o_lf0, lf0, (size_t)TF * 8);
    if (NL) memcpy(h + o_lpf, lpf, (size_t)TF * NL * 8);
  }
  r->d_in = ctx->dalloc(r->in_bytes, &r->in_cap);
  JB_CUDA(cudaMemcpyAsync(r->d_in, h, r->in_bytes, cudaMemcpyHostToDevice, ctx->stream));
  JB_CUDA(cudaEventRecord(ctx->ev[1], ctx->stream));
  r->h2d_bytes = (int64_t)r->in_bytes;

  Carver wk;
  const size_t w_b = wk.take((size_t)TF * M * 8), w_pitch = wk.take((size_t)TF * 8);
  const size_t w_pulse = wk.take((size_t)r->total_samples * 8), w_runs = wk.take(((size_t)TF + 2) * 4);
  const size_t w_wave = wk.take((size_t)r->total_samples * 8), w_out = wk.take((size_t)r->total_samples * 8);
  const size_t seg_state_doubles =
      cfg.stage != 0 ? (size_t)jbk::mglsa_state_doubles((int)cfg.stage) : (size_t)jbk::kMlsaSegStateDoubles;
  const size_t w_segstate = wk.take(plan.segs.size() * 2 * seg_state_doubles * 8);
  const size_t w_worst = wk.take(8);
  const size_t w_segmis = wk.take(plan.segs.size() * 8);
  const size_t w_failcnt = wk.take(8);
  const size_t fail_cap = std::max<size_t>(64, plan.segs.size() / 16);
  const size_t w_faillist = wk.take(fail_cap * 4);
  const bool postfilter = cfg.stage == 0 && cfg.beta > 0.0 && M > 2;
  const bool first_rows = postfilter || cfg.stage != 0 || cfg.coefficient_rows;
  r->rows_given = cfg.coefficient_rows;
  const size_t w_bfirst = first_rows ? wk.take(n_utt * M * 8) : 0;
  const bool need_binc = cfg.stage != 0 || !plan.segs.empty();
  const size_t w_binc = need_binc ? wk.take((size_t)TF * M * 8) : 0;
  r->d_work = ctx->dalloc(wk.off, &r->work_cap);
  uint8_t* di = (uint8_t*)r->d_in;
  uint8_t* dw = (uint8_t*)r->d_work;

  r->layout.n_utt = (int)n_utt;
  r->layout.frame_off = (const int64_t*)(di + o_foff);
  r->layout.order = (const int32_t*)(di + o_order);
  r->layout.max_frames = (int)r->max_frames;
  r->vp.nmcp = (int)M;
  r->vp.nlpf = (int)NL;
  r->vp.fperiod = fp;
  r->vp.rate = (double)cfg.rate;
  r->vp.alpha = cfg.alpha;
  r->vp.volume = cfg.volume;
  r->vp.stage = (int)cfg.stage;
  r->vp.use_log_gain = cfg.use_log_gain ? 1 : 0;
  r->lsp_beta = cfg.stage != 0 ? cfg.beta : 0.0;
  r->beta = postfilter ? cfg.beta : 0.0;
  r->vd.mcp = (const double*)(di + o_spec);
  r->vd.lf0 = (const double*)(di + o_lf0);
  r->vd.lpf = (const double*)(di + o_lpf);
  r->vd.b = (double*)(dw + w_b);
  r->vd.b_first = first_rows ? (const double*)(dw + w_bfirst) : nullptr;
  r->vd.binc = need_binc ? (double*)(dw + w_binc) : nullptr;
  r->vd.pitch = (double*)(dw + w_pitch);
  r->vd.pulse = (double*)(dw + w_pulse);
  r->vd.runs = (int32_t*)(dw + w_runs);
  r->vd.wave = (double*)(dw + w_wave);
  r->vd.out = (double*)(dw + w_out);
  r->vd.state = nullptr;
  r->vd.frame_begin = 0;
  r->vd.frame_count = -1;
  r->segs.n_seg = (int)plan.segs.size();
  r->segs.seg = (const jbk::MlsaSegment*)(di + o_segs);
  r->segs.state = plan.segs.empty() ? nullptr : (double*)(dw + w_segstate);
  r->segs.mismatch = plan.segs.empty() ? nullptr : (double*)(dw + w_segmis);
  r->segs.fail_count = plan.segs.empty() ? nullptr : (int*)(dw + w_failcnt);
  r->segs.fail_list = plan.segs.empty() ? nullptr : (int*)(dw + w_faillist);
  r->segs.fail_cap = plan.segs.empty() ? 0 : (int)fail_cap;
  r->seg_table = plan.segs;
  r->warm_frames = (int)(mlsa_warm_frames * (cfg.stage != 0 ? kMglsaWarmFactor : 1));
  r->whole.n_seg = (int)whole.size();
  r->whole.seg = (const jbk::MlsaSegment*)(di + o_whole);
  r->d_worst = (double*)(dw + w_worst);
  r->result = r->vd.wave;

  const size_t need = (size_t)r->max_frames * fp;
  if (need > ctx->noise_len) {
    std::lock_guard<std::mutex> gn(noise_mu_);
    const std::vector<double>& tbl = noise_table(need + need / 2);
    if (ctx->d_noise) JB_CUDA(cudaFree(ctx->d_noise));
    JB_CUDA(cudaMalloc((void**)&ctx->d_noise, tbl.size() * 8));
    JB_CUDA(cudaMemcpy(ctx->d_noise, tbl.data(), tbl.size() * 8, cudaMemcpyHostToDevice));
    ctx->noise_len = tbl.size();
  }
  r->vd.noise = ctx->d_noise;
  JB_CUDA(cudaEventSynchronize(ctx->ev[1]));
  return r.release();
}

void Engine::mlpg_create(const MlpgInput& in, int device_index, std::vector<double>* out,
                         std::vector<size_t>* frame_off) const {
  if (device_index < 0 || (size_t)device_index >= devices_.size()) throw Error(kErrArgument, "device index out of range");
  if (in.dim == 0 || in.windows.empty() || !in.state_off || !in.mean || !in.vari || (in.n_utt && !in.durations))
    throw Error(kErrArgument, "mlpg: dim, windows, state_off, durations, mean and vari are required");
  if (in.windows.size() > (size_t)jbk::kMaxWin) throw Error(kErrUnsupported, "more than 4 windows per stream");
  if ((in.gv_mean != nullptr) != (in.gv_vari != nullptr) || (in.gv_mean && !in.gv_switch))
    throw Error(kErrArgument, "mlpg: gv_mean, gv_vari and gv_switch come together");
  jbk::WindowSet ws{};
  ws.nwin = (int)in.windows.size();
  for (size_t i = 0; i < in.windows.size(); ++i) {
    const std::vector<double>& w = in.windows[i];
    if (w.empty() || w.size() % 2 == 0 || w.size() > (size_t)jbk::kMaxWinWidth)
      throw Error(kErrUnsupported, "window widths are odd and at most 5");
    ws.width[i] = (int)w.size();
    for (size_t k = 0; k < w.size(); ++k) ws.coef[i][k] = w[k];
  }
  const size_t U = in.n_utt, dim = in.dim, L = dim * in.windows.size();
  if (in.state_off[0] != 0) throw Error(kErrArgument, "mlpg: state_off must ascend from 0");
  const size_t S = in.state_off[U];
  std::vector<int64_t> soff(U + 1), foff(U + 1, 0);
  int64_t max_frames = 0;
  for (size_t u = 0; u < U; ++u) {
    if (in.state_off[u + 1] < in.state_off[u]) throw Error(kErrArgument, "mlpg: state_off must ascend from 0");
    int64_t t = 0;
    for (size_t i = in.state_off[u]; i < in.state_off[u + 1]; ++i) {
      if (in.durations[i] < 0) throw Error(kErrArgument, "mlpg: negative duration");
      t += in.durations[i];
    }
    foff[u + 1] = foff[u] + t;
    max_frames = std::max(max_frames, t);
  }
  for (size_t u = 0; u <= U; ++u) soff[u] = (int64_t)in.state_off[u];
  const int64_t TF = foff[U];
  frame_off->assign(U + 1, 0);
  for (size_t u = 0; u <= U; ++u) (*frame_off)[u] = (size_t)foff[u];
  out->assign((size_t)TF * dim, 0.0);
  if (TF == 0) {
    ctx(devices_[(size_t)device_index]);  // touch the device: there is no CPU path
    return;
  }
  const bool msd = in.msd != nullptr, gv = in.gv_mean != nullptr;
  DeviceCtx* c = ctx(devices_[(size_t)device_index]);
  std::lock_guard<std::mutex> g(c->mu);
  JB_CUDA(cudaSetDevice(c->device));
  Carver wk;
  const size_t o_soff = wk.take((U + 1) * 8), o_foff = wk.take((U + 1) * 8), o_dur = wk.take(S * 4);
  const size_t o_mean = wk.take(S * L * 8), o_vari = wk.take(S * L * 8), o_msd = msd ? wk.take(S * 8) : 0;
  const size_t o_gm = gv ? wk.take(U * dim * 8) : 0, o_gv = gv ? wk.take(U * dim * 8) : 0, o_gs = gv ? wk.take(S) : 0;
  const size_t in_bytes = wk.off;
  const size_t o_fstate = wk.take((size_t)TF * 4), o_ivar = wk.take(S * L * 8);
  const size_t o_left = msd ? wk.take((size_t)TF * 4) : 0, o_right = msd ? wk.take((size_t)TF * 4) : 0;
  const size_t o_vidx = msd ? wk.take((size_t)TF * 4) : 0, o_nv = msd ? wk.take(U * 4) : 0;
  const size_t o_par = wk.take((size_t)TF * dim * 8);
  const size_t o_scratch = wk.take((size_t)jbk::mlpg_scratch_planes(ws.band_width(), gv ? 1 : 0) * (size_t)TF * dim * 8);
  size_t cap = 0;
  uint8_t* d = (uint8_t*)c->dalloc(wk.off, &cap);
  std::vector<uint8_t> h(in_bytes, 0);
  memcpy(h.data() + o_soff, soff.data(), (U + 1) * 8);
  memcpy(h.data() + o_foff, foff.data(), (U + 1) * 8);
  memcpy(h.data() + o_dur, in.durations, S * 4);
  memcpy(h.data() + o_mean, in.mean, S * L * 8);
  memcpy(h.data() + o_vari, in.vari, S * L * 8);
  if (msd) memcpy(h.data() + o_msd, in.msd, S * 8);
  if (gv) {
    memcpy(h.data() + o_gm, in.gv_mean, U * dim * 8);
    memcpy(h.data() + o_gv, in.gv_vari, U * dim * 8);
    memcpy(h.data() + o_gs, in.gv_switch, S);
  }
  try {
    JB_CUDA(cudaMemcpyAsync(d, h.data(), in_bytes, cudaMemcpyHostToDevice, c->stream));
    jbk::BatchLayout bl{};
    bl.n_utt = (int)U;
    bl.state_off = (const int64_t*)(d + o_soff);
    bl.frame_off = (const int64_t*)(d + o_foff);
    bl.durations = (const int32_t*)(d + o_dur);
    bl.frame_state = (int32_t*)(d + o_fstate);
    bl.max_frames = (int)max_frames;
    bl.order = nullptr;
    jbk::StreamSet ss{};
    ss.n = 1;
    jbk::StreamDev& sd = ss.s[0];
    sd.dim = (int)dim;
    sd.nwin = ws.nwin;
    sd.is_msd = msd ? 1 : 0;
    sd.has_gv = gv ? 1 : 0;
    sd.windows = ws;
    sd.msd_threshold = in.msd_threshold;
    sd.gv_weight = in.gv_weight;
    sd.mean = (const double*)(d + o_mean);
    sd.vari = (const double*)(d + o_vari);
    sd.ivar = (double*)(d + o_ivar);
    sd.msd = msd ? (const double*)(d + o_msd) : nullptr;
    sd.gv_mean = gv ? (const double*)(d + o_gm) : nullptr;
    sd.gv_vari = gv ? (const double*)(d + o_gv) : nullptr;
    sd.gv_switch = gv ? (const uint8_t*)(d + o_gs) : nullptr;
    sd.left = msd ? (int32_t*)(d + o_left) : nullptr;
    sd.right = msd ? (int32_t*)(d + o_right) : nullptr;
    sd.vidx = msd ? (int32_t*)(d + o_vidx) : nullptr;
    sd.nvoiced = msd ? (int32_t*)(d + o_nv) : nullptr;
    sd.scratch = (double*)(d + o_scratch);
    sd.par = (double*)(d + o_par);
    JB_LAUNCH(jbk::launch_expand(bl, ss, c->stream));
    JB_LAUNCH(jbk::launch_mlpg(bl, sd, TF, (int64_t)S, c->stream));
    JB_CUDA(cudaMemcpyAsync(out->data(), sd.par, (size_t)TF * dim * 8, cudaMemcpyDeviceToHost, c->stream));
    JB_CUDA(cudaStreamSynchronize(c->stream));
  } catch (...) {
    cudaStreamSynchronize(c->stream);
    c->dfree(d, cap);
    throw;
  }
  c->dfree(d, cap);
}

double Engine::run_resident(Resident* r, BatchTimings* tm) const {
  DeviceCtx* c = r->ctx;
  std::lock_guard<std::mutex> g(c->mu);
  JB_CUDA(cudaSetDevice(c->device));
  JB_CUDA(cudaEventRecord(c->ev[2], c->stream));
  if (r->from_tracks) {  // the parameter tracks are the input: per-frame prep only
    JB_LAUNCH(jbk::launch_frame_prep(r->layout, r->vp, r->vd, r->total_frames, 1, c->stream));
    JB_LAUNCH(jbk::launch_pulse(r->layout, r->vp, r->vd, r->total_samples, c->stream));
    if (r->rows_given) {
      const size_t rows = (size_t)r->total_frames * r->vp.nmcp * 8;
      JB_CUDA(cudaMemcpyAsync(r->vd.b, r->vd.mcp, rows, cudaMemcpyDeviceToDevice, c->stream));
      JB_CUDA(cudaMemcpyAsync(const_cast<double*>(r->vd.b_first), (const uint8_t*)r->vd.mcp + rows,
                              (size_t)r->n_utt * r->vp.nmcp * 8, cudaMemcpyDeviceToDevice, c->stream));
      if (r->vd.binc) JB_LAUNCH(jbk::launch_row_increments(r->layout, r->vp, r->vd, r->total_frames, c->stream));
    } else {
      launch_spectrum_prep(r, c->stream);
    }
  } else {
    launch_mlpg_all(r, true);
  }
  JB_CUDA(cudaEventRecord(c->ev[3], c->stream));
  launch_vocoder(r, mlsa_variant, false, c->ev[4], true, mlsa_cut_tolerance);
  JB_CUDA(cudaEventRecord(c->ev[5], c->stream));
  JB_CUDA(cudaEventSynchronize(c->ev[5]));
  if (mlsa_cut_check_failed(r, mlsa_cut_tolerance)) {  // timed too: ev[5] is recorded again
    if (!mlsa_repair_segments(r, mlsa_cut_tolerance, nullptr)) mlsa_sequential_rerun(r);
    JB_CUDA(cudaEventRecord(c->ev[5], c->stream));
    JB_CUDA(cudaEventSynchronize(c->ev[5]));
  }
  float total = 0, a = 0, b = 0, m = 0;
  JB_CUDA(cudaEventElapsedTime(&total, c->ev[2], c->ev[5]));
  JB_CUDA(cudaEventElapsedTime(&a, c->ev[2], c->ev[3]));
  JB_CUDA(cudaEventElapsedTime(&b, c->ev[3], c->ev[4]));
  JB_CUDA(cudaEventElapsedTime(&m, c->ev[4], c->ev[5]));
  if (tm) {
    tm->kernels_ms = total;
    tm->mlpg_ms = a;
    tm->excitation_ms = b;
    tm->mlsa_ms = m;
    tm->total_frames = r->total_frames;
    tm->total_samples = r->total_samples;
    tm->voiced_samples = r->voiced_samples;
    tm->h2d_bytes = r->h2d_bytes;
    tm->launches = r->launches() + (r->used_fallback ? 1 : 0);
    tm->mlsa_segments = r->segs.n_seg;
    tm->mlsa_seg_frames = r->seg_frames;
    tm->mlsa_cut_worst = (r->segs.n_seg > 0 && r->h_worst) ? *r->h_worst : 0.0;
    tm->mlsa_fallback = r->used_fallback ? 1 : 0;
    tm->mlsa_repaired = r->repaired_segments;
  }
  return total;
}

void Engine::fetch_resident(Resident* r, double** out, std::vector<size_t>* out_off) const {
  DeviceCtx* c = r->ctx;
  std::lock_guard<std::mutex> g(c->mu);
  JB_CUDA(cudaSetDevice(c->device));
  double* h = (double*)pinned().alloc((size_t)r->total_samples * 8);
  cudaError_t e = cudaMemcpyAsync(h, r->result ? r->result : r->vd.wave, (size_t)r->total_samples * 8,
                                  cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) {
    pinned().release(h);
    throw Error(kErrCuda, std::string("CUDA error: ") + cudaGetErrorString(e));
  }
  out_off->assign((size_t)r->n_utt + 1, 0);
  for (int u = 0; u <= r->n_utt; ++u) (*out_off)[(size_t)u] = (size_t)(r->frame_off[(size_t)u] * r->vp.fperiod);
  *out = h;
}

void Engine::resident_device_output(Resident* r, const double** d_out, std::vector<size_t>* out_off, int* device) const {
  DeviceCtx* c = r->ctx;
  std::lock_guard<std::mutex> g(c->mu);
  JB_CUDA(cudaSetDevice(c->device));
  JB_CUDA(cudaStreamSynchronize(c->stream));
  *d_out = r->result ? r->result : r->vd.wave;
  if (device) *device = c->device;
  out_off->assign((size_t)r->n_utt + 1, 0);
  for (int u = 0; u <= r->n_utt; ++u) (*out_off)[(size_t)u] = (size_t)(r->frame_off[(size_t)u] * r->vp.fperiod);
}

std::vector<double> Engine::resident_rows(Resident* r) const {
  DeviceCtx* c = r->ctx;
  std::lock_guard<std::mutex> g(c->mu);
  JB_CUDA(cudaSetDevice(c->device));
  JB_CUDA(cudaStreamSynchronize(c->stream));
  const size_t M = (size_t)r->vp.nmcp, nf = (size_t)r->total_frames * M, nu = (size_t)r->n_utt * M;
  std::vector<double> out(nf + nu, 0.0);
  if (nf) JB_CUDA(cudaMemcpy(out.data(), r->vd.b, nf * 8, cudaMemcpyDeviceToHost));
  if (nu && r->vd.b_first) JB_CUDA(cudaMemcpy(out.data() + nf, r->vd.b_first, nu * 8, cudaMemcpyDeviceToHost));
  else if (nu)  // beta == 0, stage 0: the first rows are rows of vd.b
    for (int u = 0; u < r->n_utt; ++u)
      if (r->frame_off[(size_t)u + 1] > r->frame_off[(size_t)u])
        memcpy(out.data() + nf + (size_t)u * M, out.data() + (size_t)r->frame_off[(size_t)u] * M, M * 8);
  return out;
}

void Engine::release(Resident* r) const {
  if (!r) return;
  DeviceCtx* c = r->ctx;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  c->dfree(r->d_in, r->in_cap);
  c->dfree(r->d_work, r->work_cap);
  if (r->d_chunk) c->dfree(r->d_chunk, r->chunk_cap);
  delete r;
}

// Engine::synthesize for a batch (engine.rs:294-296 per utterance). Contiguous utterance ranges,
// balanced by frame count, go to the engine's GPUs; no collective is involved.
void Engine::synthesize_batch(const std::vector<std::vector<std::string>>& utts, double** out,
                              std::vector<size_t>* out_off, BatchDebug* dbg, BatchTimings* tm) const {
  synthesize_batch_impl(utts, reinterpret_cast<void**>(out), out_off, dbg, tm, false);
}

void Engine::synthesize_batch_pcm16(const std::vector<std::vector<std::string>>& utts, int16_t** out,
                                    std::vector<size_t>* out_off, BatchTimings* tm) const {
  synthesize_batch_impl(utts, reinterpret_cast<void**>(out), out_off, nullptr, tm, true);
}

void Engine::synthesize_batch_impl(const std::vector<std::vector<std::string>>& utts, void** out,
                                   std::vector<size_t>* out_off, BatchDebug* dbg, BatchTimings* tm,
                                   bool pcm16) const {
  const double t0 = now_ms();
  const size_t n = utts.size();
  const int fperiod = (int)condition.fperiod;
  if (pcm16 && fperiod % 4 != 0)
    throw Error(kErrUnsupported, "16-bit PCM delivery needs fperiod % 4 == 0 (the quantiser packs four samples per store)");
  // touch the device even for an empty batch: there is no CPU path to fall back to
  std::vector<DeviceCtx*> ctxs;
  for (int d : devices_) ctxs.push_back(ctx(d));
  const size_t ndev = dbg ? 1 : std::min(ctxs.size(), std::max<size_t>(n, 1));

  // ---- frames per utterance -> output offsets. A large batch: only the label and duration part of the
  // front-end runs here (utterance_frames); the stream parameters are looked up chunk by chunk below (by the host
  // thread of each device, on its share of the cores), so that the host works on chunk k + 1 while the GPU
  // filters chunk k. Otherwise the whole front-end runs now.
  const bool pipelined = !dbg && n >= 64 * ndev && pipeline_chunks > 1;
  std::vector<UtterancePack> all_packs;
  std::vector<size_t> nframes(n, 0);
  if (pipelined) {
    if (condition.stage != 0 || voices_.global_metadata().num_streams < 3)
      build_packs(utts, 0, 0);  // raises the proper Unsupported error
    std::unique_ptr<Error> err;
    std::mutex emu;
    parallel_for(n, [&](size_t u) {
      try {
        nframes[u] = utterance_frames(voices_, condition, utts[u]);
      } catch (const Error& e) {
        std::lock_guard<std::mutex> g(emu);
        if (!err) err.reset(new Error(e));
      } catch (const std::exception& e) {
        std::lock_guard<std::mutex> g(emu);
        if (!err) err.reset(new Error(kErrArgument, e.what()));
      }
    });
    if (err) throw *err;
  } else {
    all_packs = build_packs(utts);
    for (size_t u = 0; u < n; ++u) nframes[u] = all_packs[u].nframes;
  }
  double front_ms = now_ms() - t0;  // exposed front-end time (later chunks run under the GPU)
  out_off->assign(n + 1, 0);
  for (size_t u = 0; u < n; ++u) (*out_off)[u + 1] = (*out_off)[u] + nframes[u] * (size_t)fperiod;
  const size_t total = (*out_off)[n];
  const size_t sample_bytes = pcm16 ? 2 : 8;
  void* const host_raw = pinned().alloc(total * sample_bytes);
  double* const host = pcm16 ? nullptr : (double*)host_raw;
  int16_t* const host16 = pcm16 ? (int16_t*)host_raw : nullptr;

  // contiguous utterance ranges of about equal frame count
  auto split = [&](size_t a, size_t b, size_t parts) {
    std::vector<size_t> cut(parts + 1, b);
    cut[0] = a;
    const size_t s0 = (*out_off)[a], span = (*out_off)[b] - s0;
    size_t u = a;
    for (size_t k = 1; k < parts; ++k) {
      const size_t target = s0 + span / parts * k;
      while (u < b && (*out_off)[u] < target) ++u;
      cut[k] = u;
    }
    return cut;
  };
  const std::vector<size_t> cut = split(0, n, ndev);
  std::vector<BatchTimings> tms(ndev);
  std::vector<std::unique_ptr<Error>> errs(ndev);

  struct Chunk {
    Resident* r = nullptr;
    size_t u0 = 0, u1 = 0;
    int slot = 0;
    double pack_ms = 0;
  };
  std::vector<double> front_first(ndev, 0.0);  // exposed front-end of every device's first chunk
  auto work = [&](size_t k) {
    DeviceCtx* c = ctxs[k];
    t_host_share = ndev;
    struct Restore {
      ~Restore() { t_host_share = 1; }
    } restore_share;
    std::vector<Chunk> chunks;
    auto free_chunks = [&]() {
      cudaStreamSynchronize(c->stream);
      cudaStreamSynchronize(c->stream_copy);
      for (Chunk& ch : chunks)
        if (ch.r) {
          c->dfree(ch.r->d_in, ch.r->in_cap);
          c->dfree(ch.r->d_work, ch.r->work_cap);
          delete ch.r;
          ch.r = nullptr;
        }
      c->use_slot(0);
    };
    // taken before the try block: the error path (free_chunks) touches the device context too
    std::unique_lock<std::mutex> g(c->mu, std::defer_lock);
    try {
      const size_t u0 = cut[k], u1 = cut[k + 1];
      if (u1 <= u0) return;
      g.lock();
      // chunks of this device: each is a complete pass (gather ... filter ... copy) of its own
      size_t nchunks = 1;
      if (pipelined) {
        const int64_t frames = (int64_t)(((*out_off)[u1] - (*out_off)[u0]) / (size_t)fperiod);
        const int64_t wave = (int64_t)c->sm_count * 4 * kSegWarpsPerSmsp * 6 * (kSegMinFrames / 2);
        nchunks = (size_t)std::max<int64_t>(1, std::min<int64_t>({(int64_t)pipeline_chunks, (int64_t)DeviceCtx::kMaxChunks,
                                                                  frames / std::max<int64_t>(wave, 1)}));
      }
      // Two chunks: a small first one (one filter wave's worth of frames), so that the GPU starts
      // after a short front-end and its waveform is crossing PCIe early, then the rest.
      std::vector<size_t> ccut = split(u0, u1, nchunks);
      std::vector<std::vector<double>> chunk_waves(nchunks);
      std::vector<double> fr = (!dbg && output_mode == 1 && !pcm16) ? output_fractions : std::vector<double>{1.0};
      if (fr.empty()) fr.assign(1, 1.0);
      if (nchunks == 2 && fr.size() >= 2) {
        // chunk 0 = the first wave, chunk 1 = all the others
        double sum = 0;
        for (double x : fr) sum += x;
        const size_t s0 = (*out_off)[u0], span = (*out_off)[u1] - s0;
        size_t u = u0;
        while (u < u1 && (double)((*out_off)[u] - s0) < (double)span * (fr[0] / sum)) ++u;
        ccut = {u0, u, u1};
        chunk_waves[0] = {1.0};
        chunk_waves[1].assign(fr.begin() + 1, fr.end());
      } else if (nchunks == 1) {
        chunk_waves[0] = fr;
      }

      auto copy_result = [&](Chunk& ch) {  // whole result of a chunk on the main stream, f64 or PCM
        Resident* r = ch.r;
        if (!r->total_samples) return;
        if (pcm16) {
          int16_t* dev16 = reinterpret_cast<int16_t*>(r->vd.pulse);  // the pulse train is consumed by now
          JB_LAUNCH(jbk::launch_pcm16(r->result, dev16, r->total_samples, c->stream));
          JB_CUDA(cudaMemcpyAsync(host16 + (*out_off)[ch.u0], dev16, (size_t)r->total_samples * 2,
                                  cudaMemcpyDeviceToHost, c->stream));
        } else {
          JB_CUDA(cudaMemcpyAsync(host + (*out_off)[ch.u0], r->result, (size_t)r->total_samples * 8,
                                  cudaMemcpyDeviceToHost, c->stream));
        }
      };

      auto copy_range = [&](Chunk& ch, const RepairRange& rr) {  // samples [n0, n1) of utterance rr.utt of the chunk
        Resident* r = ch.r;
        const int64_t base = r->frame_off[(size_t)rr.utt] * r->vp.fperiod;
        int64_t n0 = base + rr.n0, n1 = base + rr.n1;
        if (pcm16) {  // launch_pcm16 works on groups of four samples
          n0 = n0 / 4 * 4;
          n1 = std::min<int64_t>((n1 + 3) / 4 * 4, r->total_samples);
        }
        if (n1 <= n0) return;
        if (pcm16) {
          int16_t* dev16 = reinterpret_cast<int16_t*>(r->vd.pulse);
          JB_LAUNCH(jbk::launch_pcm16(r->result + n0, dev16 + n0, n1 - n0, c->stream));
          JB_CUDA(cudaMemcpyAsync(host16 + (*out_off)[ch.u0] + n0, dev16 + n0, (size_t)(n1 - n0) * 2,
                                  cudaMemcpyDeviceToHost, c->stream));
        } else {
          JB_CUDA(cudaMemcpyAsync(host + (*out_off)[ch.u0] + n0, r->result + n0, (size_t)(n1 - n0) * 8,
                                  cudaMemcpyDeviceToHost, c->stream));
        }
      };

      for (size_t q = 0; q < nchunks; ++q) {
        Chunk ch;
        ch.u0 = ccut[q];
        ch.u1 = ccut[q + 1];
        ch.slot = (int)q;
        if (ch.u1 <= ch.u0) continue;
        std::vector<UtterancePack> local;
        const UtterancePack* packs = nullptr;
        if (pipelined) {
          const double f0 = now_ms();
          local = build_packs(utts, ch.u0, ch.u1 - ch.u0);  // under the GPU work of the chunk before
          for (size_t u = ch.u0; u < ch.u1; ++u)
            if (local[u - ch.u0].nframes != nframes[u]) throw Error(kErrArgument, "front-end frame count changed");
          packs = local.data();
          if (q == 0) front_first[k] = now_ms() - f0;
        } else {
          packs = all_packs.data() + ch.u0;
        }
        c->use_slot(ch.slot);
        const double p0 = now_ms();
        ch.r = make_resident(*this, c, packs, ch.u1 - ch.u0, false, condition, chunk_waves[q]);
        ch.pack_ms = now_ms() - p0;
        chunks.push_back(ch);
        Resident* r = ch.r;
        JB_CUDA(cudaEventRecord(c->ev[2], c->stream));
        launch_mlpg_all(r, true);
        JB_CUDA(cudaEventRecord(c->ev[3], c->stream));
        WaveCopy wc;
        if (pcm16) {
          wc.host16 = host16 + (*out_off)[ch.u0];
          wc.dev16 = reinterpret_cast<int16_t*>(r->vd.pulse);
        } else {
          wc.host = host + (*out_off)[ch.u0];
        }
        wc.copy_stream = c->stream_copy;
        wc.done = c->ev[6];
        wc.direct = output_mode == 2;
        launch_vocoder(r, mlsa_variant, dbg && dbg->stop_after_excitation, c->ev[4], true, mlsa_cut_tolerance,
                       (dbg || output_mode == 0) ? nullptr : &wc);
        JB_CUDA(cudaEventRecord(c->ev[5], c->stream));
        if (!r->copied_by_waves) {
          copy_result(chunks.back());
          JB_CUDA(cudaEventRecord(c->ev[6], c->stream));
        }
        if (dbg) {
          const jbk::StreamSet& ss = r->streams;
          const size_t TF = (size_t)r->total_frames;
          if (dbg->want_parameters) {
            dbg->mcp.resize(TF * ss.s[0].dim);
            dbg->lf0.resize(TF * ss.s[1].dim);
            dbg->lpf.resize(TF * ss.s[2].dim);
            JB_CUDA(cudaMemcpyAsync(dbg->mcp.data(), ss.s[0].par, dbg->mcp.size() * 8, cudaMemcpyDeviceToHost, c->stream));
            JB_CUDA(cudaMemcpyAsync(dbg->lf0.data(), ss.s[1].par, dbg->lf0.size() * 8, cudaMemcpyDeviceToHost, c->stream));
            JB_CUDA(cudaMemcpyAsync(dbg->lpf.data(), ss.s[2].par, dbg->lpf.size() * 8, cudaMemcpyDeviceToHost, c->stream));
          }
          if (dbg->want_pulse) {
            dbg->pulse.resize((size_t)r->total_samples);
            JB_CUDA(cudaMemcpyAsync(dbg->pulse.data(), r->vd.pulse, dbg->pulse.size() * 8, cudaMemcpyDeviceToHost, c->stream));
          }
          dbg->frame_off = r->frame_off;
        }
      }
      // All kernels done; the copies of the last waves are still crossing PCIe on the copy stream. A failed cut
      // is repaired now, under that copy tail: the repair kernel runs on the main stream, and only the final
      // copy of the repaired sample ranges waits for the bulk copies (same destination, so it must come last).
      JB_CUDA(cudaStreamSynchronize(c->stream));
      if (pcm16) JB_CUDA(cudaStreamSynchronize(c->stream_copy));  // the repair's scratch is the PCM staging buffer
      for (Chunk& ch : chunks) {
        Resident* r = ch.r;
        c->use_slot(ch.slot);
        if (!mlsa_cut_check_failed(r, mlsa_cut_tolerance)) continue;
        std::vector<RepairRange> ranges;
        if (mlsa_repair_segments(r, mlsa_cut_tolerance, &ranges)) {
          JB_CUDA(cudaEventRecord(c->ev[5], c->stream));
          if (r->copied_by_waves) JB_CUDA(cudaStreamWaitEvent(c->stream, c->ev[6], 0));
          for (const RepairRange& rr : ranges) copy_range(ch, rr);
        } else {  // the whole chunk without cuts
          mlsa_sequential_rerun(r);
          JB_CUDA(cudaEventRecord(c->ev[5], c->stream));
          if (r->copied_by_waves) JB_CUDA(cudaStreamWaitEvent(c->stream, c->ev[6], 0));
          copy_result(ch);
        }
        JB_CUDA(cudaEventRecord(c->ev[6], c->stream));
      }
      JB_CUDA(cudaStreamSynchronize(c->stream));
      JB_CUDA(cudaStreamSynchronize(c->stream_copy));

      BatchTimings& t = tms[k];
      for (Chunk& ch : chunks) {
        Resident* r = ch.r;
        c->use_slot(ch.slot);
        t.mlsa_segments += r->segs.n_seg;
        t.mlsa_seg_frames += r->seg_frames;
        const double worst = r->segs.n_seg > 0 && r->h_worst && !(dbg && dbg->stop_after_excitation) ? *r->h_worst : 0.0;
        t.mlsa_cut_worst = std::max(t.mlsa_cut_worst, worst);
        t.mlsa_fallback += r->used_fallback ? 1 : 0;
        t.mlsa_repaired += r->repaired_segments;
        float f = 0;
        cudaEvent_t kend = (r->kernels_end && !r->used_fallback && !r->repaired_segments) ? r->kernels_end : c->ev[5];
        JB_CUDA(cudaEventElapsedTime(&f, c->ev[0], c->ev[1])); t.h2d_ms += f;
        JB_CUDA(cudaEventElapsedTime(&f, c->ev[2], c->ev[3])); t.mlpg_ms += f;
        JB_CUDA(cudaEventElapsedTime(&f, c->ev[3], c->ev[4])); t.excitation_ms += f;
        JB_CUDA(cudaEventElapsedTime(&f, c->ev[4], kend)); t.mlsa_ms += f;
        if (ch.slot == chunks.front().slot) t.pack_ms = ch.pack_ms;  // later chunks pack under the GPU
        t.h2d_bytes += r->h2d_bytes;
        t.d2h_bytes += r->total_samples * (int64_t)sample_bytes;
        t.total_frames += r->total_frames;
        t.total_samples += r->total_samples;
        t.voiced_samples += r->voiced_samples;
        t.launches += r->launches() + (r->used_fallback ? 1 : 0);
      }
      if (!chunks.empty()) {
        // span of all kernels (first chunk's start to the last chunk's last kernel), then the copy tail
        Chunk& a = chunks.front();
        Chunk& z = chunks.back();
        cudaEvent_t* eva = c->evsets[a.slot];
        cudaEvent_t* evz = c->evsets[z.slot];
        cudaEvent_t kend = (z.r->kernels_end && !z.r->used_fallback && !z.r->repaired_segments) ? z.r->kernels_end : evz[5];
        float f = 0;
        JB_CUDA(cudaEventElapsedTime(&f, eva[2], kend)); t.kernels_ms = f;
        JB_CUDA(cudaEventElapsedTime(&f, kend, evz[6])); t.d2h_ms = f > 0 ? f : 0;
      }
      free_chunks();
    } catch (const Error& e) {
      free_chunks();
      errs[k].reset(new Error(e));
    } catch (const std::exception& e) {
      free_chunks();
      errs[k].reset(new Error(kErrCuda, e.what()));
    }
  };
  if (ndev == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (size_t k = 0; k < ndev; ++k) th.emplace_back(work, k);
    for (auto& t : th) t.join();
  }
  for (auto& e : errs)
    if (e) {
      pinned().release(host_raw);
      throw *e;
    }
  front_ms += *std::max_element(front_first.begin(), front_first.end());
  if (tm) {
    *tm = BatchTimings();
    tm->host_frontend_ms = front_ms;
    for (const auto& t : tms) {
      tm->pack_ms = std::max(tm->pack_ms, t.pack_ms);
      tm->h2d_ms = std::max(tm->h2d_ms, t.h2d_ms);
      tm->kernels_ms = std::max(tm->kernels_ms, t.kernels_ms);
      tm->mlpg_ms = std::max(tm->mlpg_ms, t.mlpg_ms);
      tm->excitation_ms = std::max(tm->excitation_ms, t.excitation_ms);
      tm->mlsa_ms = std::max(tm->mlsa_ms, t.mlsa_ms);
      tm->d2h_ms = std::max(tm->d2h_ms, t.d2h_ms);
      tm->h2d_bytes += t.h2d_bytes;
      tm->d2h_bytes += t.d2h_bytes;
      tm->total_frames += t.total_frames;
      tm->total_samples += t.total_samples;
      tm->voiced_samples += t.voiced_samples;
      tm->launches += t.launches;
      tm->mlsa_segments += t.mlsa_segments;
      tm->mlsa_seg_frames += t.mlsa_seg_frames;
      tm->mlsa_cut_worst = std::max(tm->mlsa_cut_worst, t.mlsa_cut_worst);
      tm->mlsa_fallback += t.mlsa_fallback;
      tm->mlsa_repaired += t.mlsa_repaired;
    }
  }
  *out = host_raw;
}

// ---------------------------------------------------------------- streaming
Stream::Stream(const Engine* e, const std::vector<std::vector<std::string>>& utts, int device)
    : e_(e), r_(nullptr) {
  std::vector<UtterancePack> packs = e->build_packs(utts);
  // `device` is an index into the engine's device list (jb_engine_set_devices), like everywhere in the C ABI
  if (device < 0 || (size_t)device >= e->devices().size()) throw Error(kErrArgument, "device index out of range");
  DeviceCtx* c = e->ctx(e->devices()[(size_t)device]);
  std::lock_guard<std::mutex> g(c->mu);
  r_ = Engine::make_resident(*e, c, packs.data(), packs.size(), true, e->condition, {});
  try {
    // MLPG for the whole utterance up front, like Engine::generator (engine.rs:333-357)
    launch_mlpg_all(r_, false);
    if (r_->n_utt)
      JB_CUDA(cudaMemsetAsync(r_->vd.state, 0, (size_t)r_->n_utt * jbk::kVocStateDoubles * 8, c->stream));
    JB_CUDA(cudaStreamSynchronize(c->stream));
  } catch (...) {
    cudaStreamSynchronize(c->stream);
    c->dfree(r_->d_in, r_->in_cap);
    c->dfree(r_->d_work, r_->work_cap);
    delete r_;
    r_ = nullptr;
    throw;
  }
}

Stream::~Stream() { e_->release(r_); }
size_t Stream::fperiod() const { return (size_t)r_->vp.fperiod; }
size_t Stream::n_utt() const { return (size_t)r_->n_utt; }
const std::vector<int64_t>& Stream::frames() const { return r_->frame_off; }

double Stream::step(size_t n_frames, double* out, size_t* written) {
  DeviceCtx* c = r_->ctx;
  std::lock_guard<std::mutex> g(c->mu);
  JB_CUDA(cudaSetDevice(c->device));
  const size_t fp = (size_t)r_->vp.fperiod, n = (size_t)r_->n_utt;
  const size_t chunk = n_frames * fp;
  if (n == 0 || n_frames == 0) return 0.0;
  if (r_->chunk_cap < n * chunk * 8) {
    if (r_->d_chunk) c->dfree(r_->d_chunk, r_->chunk_cap);
    r_->d_chunk = (double*)c->dalloc(n * chunk * 8, &r_->chunk_cap);
  }
  r_->vd.frame_begin = (int)next_;
  r_->vd.frame_count = (int)n_frames;
  JB_CUDA(cudaEventRecord(c->ev[2], c->stream));
  launch_vocoder(r_, e_->mlsa_variant, false, nullptr, false, e_->mlsa_cut_tolerance);
  JB_LAUNCH(jbk::launch_gather_chunk(r_->layout, r_->vp, r_->vd, r_->d_chunk, c->stream));
  JB_CUDA(cudaEventRecord(c->ev[5], c->stream));
  JB_CUDA(cudaMemcpyAsync(out, r_->d_chunk, n * chunk * 8, cudaMemcpyDeviceToHost, c->stream));
  JB_CUDA(cudaStreamSynchronize(c->stream));
  for (size_t u = 0; u < n; ++u) {
    const int64_t T = r_->frame_off[u + 1] - r_->frame_off[u];
    int64_t left = T - (int64_t)next_;
    if (left < 0) left = 0;
    written[u] = (size_t)std::min<int64_t>(left, (int64_t)n_frames) * fp;
  }
  next_ += n_frames;
  float ms = 0;
  JB_CUDA(cudaEventElapsedTime(&ms, c->ev[2], c->ev[5]));
  return ms;
}
}  // namespace jb
