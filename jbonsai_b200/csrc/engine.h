// C++ host Engine: the drop-in for jbonsai's `Engine` (src/engine.rs:247-366) whose
// MlpgAdjust::create + SpeechGenerator::generate_all run as CUDA kernels on B200.
// Host side = label parsing, tree lookup, durations and the batching layer that packs
// state-level means/variances into device buffers (north star); there is no CPU fallback:
// every synthesis call fails with kErrCuda when no device is usable.
#pragma once

#include <atomic>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "host/frontend.h"

namespace jb {

class DeviceCtx;  // one per GPU (engine.cc)

// Optional per-call taps / counters used by tests and bench.py.
struct BatchDebug {
  bool want_parameters = false;  // copy MLPG output back: mcp [TF][35], lf0 [TF], lpf [TF][31]
  bool want_pulse = false;       // copy the pulse train back: [TS]
  bool stop_after_excitation = false;  // wave = excitation*gain (no MLSA)
  std::vector<double> mcp, lf0, lpf, pulse;
  std::vector<int64_t> frame_off;  // [n_utt+1]
};

struct BatchTimings {
  double host_frontend_ms = 0;  // labels -> packs
  double pack_ms = 0;
  double h2d_ms = 0, kernels_ms = 0, d2h_ms = 0;  // CUDA-event times on the device stream
  double mlpg_ms = 0, excitation_ms = 0, mlsa_ms = 0;
  int64_t h2d_bytes = 0, d2h_bytes = 0;
  int64_t total_frames = 0, total_samples = 0, voiced_samples = 0;
  int launches = 0;
  // MLSA variant 3 (time-segmented filter): segments, frames filtered including warm-up, the
  // largest state mismatch found at a cut, and whether the sequential kernel had to redo the batch
  int64_t mlsa_segments = 0, mlsa_seg_frames = 0;
  double mlsa_cut_worst = 0;
  int mlsa_fallback = 0;
  int mlsa_repaired = 0;  // segments re-filtered with a longer warm-up after a failed cut check
};

class Engine {
 public:
  static std::unique_ptr<Engine> load(const std::vector<std::string>& paths);
  static std::unique_ptr<Engine> load_from_bytes(const std::vector<std::vector<uint8_t>>& bufs);
  ~Engine();

  Condition condition;  // public like the reference's `engine.condition`
  const VoiceSet& voices() const { return voices_; }

  // GPUs used by synthesize*/streams. Default: device 0 only.
  void set_devices(const std::vector<int>& devices);
  const std::vector<int>& devices() const { return devices_; }
  int mlsa_variant = -1;  // -1 = default kernel (3: time-segmented; streams use 2)
  // Device-side parameter gather (default): leaf pdf tables stay resident in HBM, the host sends
  // row indices and k_gather_states interpolates. false = the host packs interpolated vectors.
  bool device_gather = true;
  // Variant 3: frames a segment filters ahead of its own range (from a zero state).
  int mlsa_warm_frames = 40;
  // Variant 3: bound on the waveform error a cut may cause. The device compares, for every cut, the state the
  // segment reached after its warm-up with the true one (the end state of the segment before it) and scales
  // the largest difference d to what it can do to the output: a tap error d moves a stage output by at most
  // d * sum_k |c_k| (the coefficients of the frame at the cut) and the waveform by about as much, times the
  // volume. A cut whose bound exceeds this value is repaired (Engine::mlsa_repair_segments); half the 1e-6 bar
  // of the north star. Two correct trajectories with different rounding histories already differ by up to
  // ~3e-8 in the delay-line taps (measured floor: nitech 4e-9, tohoku-f01 3e-8), i.e. ~2e-7 in this measure.
  // Stage != 0 (MGLSA): the state mismatch relative to the state's scale is held to 1e-9 instead.
  double mlsa_cut_tolerance = 5e-7;
  // Variant 3, stage 0: warm-up of a cut whose last `mlsa_warm_loud` frames have a mean log gain (coefficient 0 of
  // the filter rows, estimated on the host from the static means of the states' spectrum pdfs) above
  // `mlsa_loud_threshold`; every other cut keeps mlsa_warm_frames. The filter forgets its start twice as fast
  // behind speech as in a pause (tools/decay_predictor.py: output error after W frames from a zero state, 1542
  // cuts of nitech: mean log gain of the last 20 frames > 1 -> 4.4e-11 at W = 20 on every one of them, 74 % of
  // all cuts; the others: up to 5.3e-5 at W = 20, 3.5e-10 at W = 32; tohoku-f01: 1.3e-10 at W = 20 everywhere).
  // The planner places the cuts of an utterance so that its segments still filter equally many frames. 0 = off.
  // Every cut is still checked on the device and repaired if need be; a device that had to repair more than
  // 0.5 % of the cuts of two batches planned this way turns the rule off for itself.
  int mlsa_warm_loud = 24;
  // Threshold 2.0: on the eight 1024-utterance batches of the scaling run (tools/loud_sweep.py, B200) thresholds
  // 1.5-2.5 leave no cut to repair with either voice, 1.0 one or two cuts in three of the eight nitech batches
  // (windows that are half pause), 3.0 moves enough cuts into pauses that one of them fails even with 40 frames.
  double mlsa_loud_threshold = 2.0;
  // How synthesize_batch brings the waveform to the host: 0 = one copy after the filter, 1 = several
  // filter waves (output_waves), each copied under the next one (default), 2 = the filter stores
  // straight into the pinned output buffer (zero-copy; measured slower: SM stores reach ~34 GB/s
  // over PCIe against ~52 GB/s for the copy engine).
  int output_mode = 1;
  // Output mode 1: shares of the batch (by frames) filtered by successive launches; fewer when the
  // batch is small. More waves start the PCIe transfer earlier but shorten the segments, so the
  // warm-up share of the filter grows (measured e2e for 1024 x 5 s with equal waves: 1 wave 86 ms,
  // 2: 73, 3: 70.8, 4: 70.3, 6: 78). A filter launch over a share f costs ~(18.3 f + 5) ms and its copy
  // 38 f ms on a B200 behind PCIe 5: a mid-sized first wave, a large second one and a small last one
  // (whose copy nothing hides) finish earliest.
  static std::vector<double> default_output_fractions() { return {0.35, 0.40, 0.25}; }
  std::vector<double> output_fractions = default_output_fractions();
  // Sub-batches of one synthesize_batch call on one device (large batches only): the stream-parameter
  // look-ups of chunk k + 1 run on the host while the GPU works on chunk k, and chunk k's waveform
  // crosses PCIe under chunk k + 1's kernels. 1 = the whole batch in one pass; 2 = the first output share as a
  // sub-batch of its own, the other shares as waves of the second; 3 and 4 = equal sub-batches of one wave each.
  // Measured e2e, 1024 x 5 s on one B200 (tools/e2e_probe.py, p50 of 8 calls): 1 -> 66 ms, 2 -> 60.2, 3 -> 57.8,
  // 4 -> 62.9 (every sub-batch pays the warm-up of a chip-filling set of segments over fewer frames).
  int pipeline_chunks = 3;

  // Engine::synthesize for a batch: utterance u = lines[line_off[u] .. line_off[u+1]).
  // `out` is a pinned host buffer owned by the engine's pool: release with free_output().
  // out_off has n_utt+1 entries (sample offsets).
  void synthesize_batch(const std::vector<std::vector<std::string>>& utts, double** out,
                        std::vector<size_t>* out_off, BatchDebug* dbg = nullptr,
                        BatchTimings* tm = nullptr) const;
  // Same synthesis, delivered as 16-bit PCM (clamp, then truncate toward zero: what the reference's
  // examples do with the Vec<f64>, examples/genji/main.rs:34-35): a quarter of the bytes cross PCIe.
  void synthesize_batch_pcm16(const std::vector<std::vector<std::string>>& utts, int16_t** out,
                              std::vector<size_t>* out_off, BatchTimings* tm = nullptr) const;

  // Hot path only: packs already built (bench "value": inputs resident on the device is handled
  // inside by `resident` handles, see prepare/run below).
  struct Resident;  // a batch whose packed inputs already sit in HBM
  Resident* prepare(const std::vector<std::vector<std::string>>& utts, int device_index) const;
  // The vocoder seam (Vocoder::new, vocoder/mod.rs:45-70, driven frame by frame as SpeechGenerator does,
  // speech.rs:65-96) for n_utt independent parameter tracks: a resident batch whose run() skips MLPG.
  // frame_off[n_utt + 1]; spectrum [frames][nmcp] (mel-cepstrum, or LSP with the gain in column 0 when
  // stage > 0), lf0 [frames] (NODATA = -1e10 in unvoiced frames), lpf [frames][nlpf] (nlpf == 0: no LPF
  // stream, excitation.rs:87-99). Throws kErrArgument for an even nlpf > 0 (speech.rs:38-40).
  struct VocoderConfig {
    size_t nmcp = 0, nlpf = 0, stage = 0;
    bool use_log_gain = false;
    size_t rate = 48000;
    double alpha = 0.0, beta = 0.0, volume = 1.0;
    size_t fperiod = 240;
    // test seam: `spectrum` already holds filter coefficient rows, [frames][nmcp] followed by the
    // [n_utt][nmcp] rows the first frames start from; the spectrum transform is skipped
    bool coefficient_rows = false;
  };
  // Filter coefficient rows of a resident batch after run(): [frames][nmcp] then [n_utt][nmcp] first rows.
  std::vector<double> resident_rows(Resident* r) const;
  Resident* prepare_tracks(const VocoderConfig& cfg, size_t n_utt, const size_t* frame_off, const double* spectrum,
                           const double* lf0, const double* lpf, int device_index) const;
  // Runs MLPG + vocoder kernels on the resident batch; waveform stays on the device.
  // Returns device time in ms (CUDA events on the launching stream).
  double run_resident(Resident* r, BatchTimings* tm = nullptr) const;
  // Copies the resident batch's waveform to a pinned host buffer (pool-owned).
  void fetch_resident(Resident* r, double** out, std::vector<size_t>* out_off) const;
  void release(Resident* r) const;

  // Device-resident delivery: the waveform of the last run_resident stays in HBM. *d_out: its device pointer
  // (out_off[n_utt] doubles, on CUDA device *device), valid until the next run_resident / release of `r`.
  void resident_device_output(Resident* r, const double** d_out, std::vector<size_t>* out_off, int* device) const;

  // The MLPG seam (MlpgAdjust::new + create, src/mlpg_adjust/mod.rs:31-95) for one stream of n_utt utterances:
  // state-level parameters in, [frames][dim] trajectories out (NODATA = -1e10 in masked frames of an MSD stream).
  struct MlpgInput {
    size_t n_utt = 0, dim = 0;
    const size_t* state_off = nullptr;  // [n_utt + 1]
    const int32_t* durations = nullptr; // [states] frames per state (>= 0)
    std::vector<std::vector<double>> windows;  // Windows (window.rs:4-56): coefficient lists, static window first
    const double* mean = nullptr;       // [states][nwin * dim], index dim * w + d (model/mod.rs:62)
    const double* vari = nullptr;
    const double* msd = nullptr;        // [states] or nullptr (not an MSD stream)
    double msd_threshold = 0.5;
    const double* gv_mean = nullptr;    // [n_utt][dim] or nullptr (no GV)
    const double* gv_vari = nullptr;
    const uint8_t* gv_switch = nullptr; // [states]
    double gv_weight = 1.0;
  };
  // out: [frames][dim] (resized); frame_off: [n_utt + 1]
  void mlpg_create(const MlpgInput& in, int device_index, std::vector<double>* out, std::vector<size_t>* frame_off) const;

  static void free_output(void* p);
  static void* alloc_output(size_t bytes);  // pinned host memory from the same pool (jb_alloc_pinned)
  // Host threads the front-end and the batching layer use (0 = all cores; also JB_HOST_THREADS).
  static void set_host_threads(int n);
  // Frees the cached pinned host buffers of the process-wide pool; returns the bytes released.
  static size_t trim_pools();
  // Test hook (host only, no device): the time segments the planner cuts for utterances of the given frame counts -
  // `gain` (per frame, all utterances concatenated) stands for the log-gain estimate of the warm-up by loudness,
  // nullptr = one warm-up for every cut. Rows of `out`: utt, begin, end, warm, prev.
  static std::vector<int32_t> plan_segments_debug(const std::vector<int64_t>& frames, int warm, int warm_loud,
                                                  double threshold, const double* gain, int64_t target);

  std::vector<double> noise_prefix(size_t n) const;  // first n Gaussians (host table)

 private:
  void synthesize_batch_impl(const std::vector<std::vector<std::string>>& utts, void** out,
                             std::vector<size_t>* out_off, BatchDebug* dbg, BatchTimings* tm, bool pcm16) const;
  friend class DeviceCtx;
  friend class Stream;
  explicit Engine(VoiceSet v);
  VoiceSet voices_;
  std::vector<StreamTable> tables_;  // flattened leaf pdfs per stream (device gather)
  // log gain (coefficient 0 after mc2b) of the static mean of every spectrum pdf row, for the alpha it was built
  // with: the loudness estimate of the segment planner (mlsa_warm_loud)
  mutable std::mutex row_gain_mu_;
  mutable std::vector<double> row_gain_;
  mutable double row_gain_alpha_ = -1e300;
  const std::vector<double>& row_gain(double alpha) const;
  std::vector<int> devices_{0};
  mutable std::mutex ctx_mu_;
  mutable std::map<int, std::unique_ptr<DeviceCtx>> ctx_;
  DeviceCtx* ctx(int device) const;
  // global white-noise sequence (excitation.rs:178-237; utterance independent)
  mutable std::mutex noise_mu_;
  mutable std::vector<double> noise_;
  mutable uint64_t rng_next_ = 1;
  const std::vector<double>& noise_table(size_t n) const;
  std::vector<UtterancePack> build_packs(const std::vector<std::vector<std::string>>& utts, size_t first = 0,
                                         size_t count = (size_t)-1) const;
  static Resident* make_resident(const Engine& e, DeviceCtx* ctx, const UtterancePack* packs,
                                 size_t n_utt, bool streaming, const Condition& cond,
                                 const std::vector<double>& wave_weights);
};

// SpeechGenerator (src/speech.rs) for many utterances at once: MLPG up front, then the
// vocoder advanced `n_frames` at a time with its state persisted in device memory.
class Stream {
 public:
  Stream(const Engine* e, const std::vector<std::vector<std::string>>& utts, int device);
  ~Stream();
  size_t fperiod() const;
  size_t n_utt() const;
  // Generates up to n_frames frames for every utterance. out: [n_utt][n_frames*fperiod] (host),
  // written[u] = samples produced for utterance u (0 at the end). Returns device ms.
  double step(size_t n_frames, double* out, size_t* written);
  const std::vector<int64_t>& frames() const;      // total frames per utterance
  size_t synthesized_frames() const { return next_; }

 private:
  const Engine* e_;
  Engine::Resident* r_;
  size_t next_ = 0;
};

}  // namespace jb
