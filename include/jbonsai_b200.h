/* jbonsai_b200 — C ABI of the B200-native hot path of jbonsai (MLPG solve + mel-cepstral vocoder).
 *
 * This is the drop-in boundary: every entry point replaces one item of the reference's public
 * Rust API (jbonsai v0.4.2; the reference has no FFI of its own). A Rust `Engine`
 * would bind these 1:1. Plain pointers and sizes only; all functions return
 * 0 on success or a negative jb_status; jb_last_error() gives the message for the calling thread.
 * There is NO CPU fallback: synthesis entry points return JB_ERR_CUDA when no GPU is usable.
 */
#ifndef JBONSAI_B200_H_
#define JBONSAI_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum jb_status {
  JB_OK = 0,
  JB_ERR_EMPTY_VOICE = -1,   /* ModelError::EmptyVoice        src/model/mod.rs:31-46 */
  JB_ERR_METADATA = -2,      /* ModelError::MetadataError */
  JB_ERR_IO = -3,            /* ModelError::Io */
  JB_ERR_PARSE = -4,         /* ModelError::ParserError */
  JB_ERR_OPTION = -5,        /* EngineError::ParseOptionError src/engine.rs:16-28 */
  JB_ERR_LABEL_PARSE = -6,   /* LabelError::JLabelParse       src/label.rs:7-19 */
  JB_ERR_LABEL_MISSING = -7, /* LabelError::MissingLabel */
  JB_ERR_LABEL_FLOAT = -8,   /* LabelError::FloatParse */
  JB_ERR_LABEL_LENGTH = -9,  /* LabelError::LengthMismatch */
  JB_ERR_WEIGHT = -10,       /* WeightError                   src/model/interporation_weight.rs:5-11 */
  JB_ERR_INDEX = -11,        /* tree/pdf index not found (reference: todo!() panic, voice/model.rs:78) */
  JB_ERR_ARGUMENT = -12,     /* bad argument (reference: panic!, e.g. src/speech.rs:32-40,70) */
  JB_ERR_CUDA = -13,         /* CUDA failure or no device */
  JB_ERR_UNSUPPORTED = -14   /* outside the built scope (stage != 0, beta > 0, no LPF stream) */
} jb_status;

typedef struct jb_engine jb_engine;
typedef struct jb_stream jb_stream;

const char* jb_last_error(void);
const char* jb_version(void);
int jb_device_count(void); /* number of usable CUDA devices, 0 if none */

/* Roofline denominators measured live on `device` (no reference equivalent; bench.py): FP64 FMA
 * rate of a DFMA micro-kernel in TFLOP/s and read+write bandwidth of a plain copy in GB/s. */
int jb_device_measure_peaks(int device, double* fp64_fma_tflops, double* hbm_copy_gbs);

/* Parity tap: the pitch period the vocoder derives from a log-F0 value, rate / exp(clamp(lf0)) with 0 for
 * NODATA (src/vocoder/mod.rs:73-77), computed by the kernel the engine runs (device `exp`); tools/exp_sweep.py
 * compares it with the host libm over the whole lf0 range. */
int jb_device_pitch_period(int device, const double* lf0, size_t n, double rate, double* out);

/* Engine::load (src/engine.rs:257) / Engine::load_from_bytes (src/engine.rs:263). */
int jb_engine_load(const char* const* paths, size_t n_paths, jb_engine** out);
int jb_engine_load_from_bytes(const uint8_t* const* bufs, const size_t* lens, size_t n,
                              jb_engine** out);
void jb_engine_free(jb_engine* e);

/* `engine.condition` (src/engine.rs:32-243). key is one of: "speed", "volume" (dB), "alpha",
 * "beta", "additional_half_tone", "msd_threshold" (idx = stream), "gv_weight" (idx = stream),
 * "phoneme_alignment", "sampling_frequency", "fperiod", "stage", "use_log_gain".
 * Setters clamp exactly like the reference's. */
int jb_engine_set_condition(jb_engine* e, const char* key, size_t idx, double value);
int jb_engine_get_condition(const jb_engine* e, const char* key, size_t idx, double* value);
/* InterporationWeight::{set_duration,set_parameter,set_gv} (interporation_weight.rs:79-113);
 * kind: 0 duration, 1 parameter (stream), 2 gv (stream). */
int jb_engine_set_interpolation_weight(jb_engine* e, int kind, size_t stream, const double* w,
                                       size_t n);

/* CUDA devices used by this engine (default: device 0). A batch is sharded over them by
 * contiguous utterance ranges balanced by frame count (no collective). */
int jb_engine_set_devices(jb_engine* e, const int* devices, size_t n);
/* MLSA kernel variant (-1 = default = 3: time-segmented; 2: one warp per utterance, sequential;
 * 0: reference operation order). For A/B measurements only. */
int jb_engine_set_mlsa_variant(jb_engine* e, int variant);
/* Where VoiceSet::weighted (src/model/voice_set.rs:80-96) runs. 1 (default): the leaf pdf tables
 * stay resident in HBM, the host sends one row index per (state, voice) and a kernel interpolates;
 * 0: the host interpolates and copies the state-level vectors (~0.5 MB per 5 s utterance). Results
 * are bit-identical. */
int jb_engine_set_device_gather(jb_engine* e, int enabled);
/* Variant 3 cuts utterances into time segments that are filtered concurrently; each segment first
 * filters `warm_frames` frames ahead of its own range from a zero state (default 40). The states on
 * both sides of every cut are compared on the device and the largest difference d of a cut is scaled to a
 * bound on the waveform error, d * max(1, sum_k |c_k| of the frame at the cut) * volume. A cut whose bound
 * exceeds `cut_tolerance` (default 5e-7, half the 1e-6 bar) is repaired: the head of the segment behind it is
 * filtered again from the exact state; if that does not settle it the sequential kernel redoes the batch. */
int jb_engine_set_mlsa_segmenting(jb_engine* e, int warm_frames, double cut_tolerance);
/* Warm-up by loudness (variant 3, stage 0): a cut whose last `loud_frames` frames have a mean log gain (coefficient 0
 * of the filter rows) above `threshold` filters only `loud_frames` frames ahead (default 24, threshold 2.0); the
 * others keep `warm_frames` (the filter forgets its start twice as fast behind speech as in a pause). The gain is
 * estimated on the host from the state-level spectrum means; the cuts of an utterance are placed so that all its
 * segments still filter equally many frames. loud_frames = 0 turns the rule off. Cuts are checked and repaired as
 * above whatever the rule says. */
int jb_engine_set_mlsa_loud_warm(jb_engine* e, int loud_frames, double threshold);
/* How jb_engine_synthesize_batch brings the waveform (8 B per sample, the Vec<f64> of
 * engine.rs:294) to the host. 1 (default): every filter launch's part of the batch is copied under the
 * next launch (sub-batches: jb_engine_set_pipeline_chunks; 10 + w selects w equal launches instead); 2: the filter kernel stores straight into
 * the pinned, device-visible output buffer (measured slower); 0: one copy after the filter. Same
 * samples (to rounding of the segment cuts) either way. */
int jb_engine_set_output_mode(jb_engine* e, int mode);
/* Output mode 1 with explicit shares of the batch (by frames) for the successive filter launches
 * (default 0.35, 0.40, 0.25). */
int jb_engine_set_output_fractions(jb_engine* e, const double* fractions, size_t n);
/* Sub-batches of one jb_engine_synthesize_batch call per device (default 3, used for batches of
 * 64+ utterances per device that fill the chip more than once): the host looks up the stream parameters of
 * chunk k + 1 while the GPU works on chunk k, and chunk k's waveform crosses PCIe under chunk k + 1's kernels.
 * 1 = the whole batch in one pass (three copy-overlapped filter launches, jb_engine_set_output_fractions);
 * 2 = the first of those launches as a sub-batch of its own; 3, 4 = equal sub-batches. Same samples. */
int jb_engine_set_pipeline_chunks(jb_engine* e, int chunks);

/* Engine::synthesize (src/engine.rs:294-296): labels -> raw f64 samples (not clipped).
 * *out is owned by the library: release with jb_free(). Empty input -> *out_len = 0. */
int jb_engine_synthesize(const jb_engine* e, const char* const* lines, size_t n_lines, double** out,
                         size_t* out_len);

/* Batch of utterances (no reference equivalent; BASELINE configs 3-5). Utterance u owns
 * lines[line_off[u] .. line_off[u+1]). out_off must hold n_utt+1 entries: utterance u's samples
 * are (*out)[out_off[u] .. out_off[u+1]), in input order. timings (optional, 20 doubles):
 * [0] host front-end ms [1] pack ms [2] H2D ms [3] kernels ms [4] D2H ms [5] MLPG ms
 * [6] excitation ms [7] MLSA ms [8] H2D bytes [9] D2H bytes [10] frames [11] samples
 * [12] voiced samples [13] kernel launches [14] MLSA time segments [15] frames filtered including
 * warm-up [16] largest waveform-error bound of a cut in the first pass [17] passes redone without cuts
 * by the sequential kernel [18] segments whose head was re-filtered from the exact state after their cut
 * check failed. */
int jb_engine_synthesize_batch(const jb_engine* e, const char* const* lines, const size_t* line_off,
                               size_t n_utt, double** out, size_t* out_off, double* timings);
/* The same synthesis delivered as 16-bit PCM: every f64 sample x becomes
 * (int16_t) min(max(x, -32768), 32767), truncated toward zero — what the reference's examples do with
 * the Vec<f64> before writing a WAV file (examples/genji/main.rs:34-35, examples/is-bonsai/main.rs:47-48).
 * The conversion runs on the device, so 2 instead of 8 bytes per sample cross PCIe. Release *out
 * with jb_free(). */
int jb_engine_synthesize_batch_pcm16(const jb_engine* e, const char* const* lines, const size_t* line_off,
                                     size_t n_utt, int16_t** out, size_t* out_off, double* timings);
void jb_free(void* p);

/* Parity taps (tests): MlpgAdjust::create output of every utterance (src/mlpg_adjust/mod.rs:51)
 * mcp [frames][35], lf0 [frames][1], lpf [frames][31]; the pulse train [samples] and, in *exc,
 * excitation x gain before the MLSA filter. frame_off: n_utt+1 entries. Buffers: jb_free(). */
int jb_engine_debug_batch(const jb_engine* e, const char* const* lines, const size_t* line_off,
                          size_t n_utt, double** mcp, double** lf0, double** lpf, double** pulse,
                          double** exc, int64_t* frame_off);
/* First n samples of the global Gaussian excitation sequence (src/vocoder/excitation.rs:197-237). */
int jb_engine_noise(const jb_engine* e, size_t n, double* out);

/* Host front-end taps (tests of the model lookup / duration goldens). */
int jb_engine_durations(const jb_engine* e, const char* const* lines, size_t n_lines,
                        int32_t* durations, size_t cap, size_t* n_states);
int jb_engine_tree_index(const jb_engine* e, size_t voice, int model /*0 dur,1 stream,2 gv*/,
                         size_t stream, size_t state_index, const char* label, long* tree_index,
                         long* pdf_index);
/* Interpolated state-level parameters: mean/vari [n_states][nwin*dim], msd [n_states]. */
int jb_engine_stream_parameters(const jb_engine* e, const char* const* lines, size_t n_lines,
                                size_t stream, double* mean, double* vari, double* msd, size_t cap,
                                size_t* n_states, size_t* width);
int jb_engine_duration_parameters(const jb_engine* e, const char* const* lines, size_t n_lines,
                                  double* mean, double* vari, size_t cap, size_t* n_states);
int jb_engine_gv_parameters(const jb_engine* e, const char* const* lines, size_t n_lines,
                            size_t stream, double* mean, double* vari, uint8_t* gv_switch,
                            size_t cap, size_t* dim, size_t* n_states, int* has_gv);

/* Resident batch: inputs packed and copied to HBM once, kernels timed alone (bench "value"). */
typedef struct jb_resident jb_resident;
int jb_resident_prepare(const jb_engine* e, const char* const* lines, const size_t* line_off,
                        size_t n_utt, int device_index, jb_resident** out);
/* The vocoder seam below the engine: Vocoder::new(nmcp, nlpf, stage, use_log_gain, rate, alpha, beta,
 * volume, fperiod) (src/vocoder/mod.rs:45-70) driven over whole parameter tracks as
 * SpeechGenerator::new(fperiod, vocoder, spectrum, lf0, lpf).generate_all() does (src/speech.rs:25-96),
 * for n_utt independent utterances at once. frame_off[n_utt + 1] (frames, ascending from 0);
 * spectrum[frames][nmcp] = mel-cepstrum rows, or line spectral pairs with the gain in column 0 when
 * stage > 0 (Stage::NonZero, vocoder/mod.rs:142-176: MGLSA filter of `stage` sections, gamma = -1/stage);
 * lf0[frames] with -1e10 (NODATA) in unvoiced frames; lpf[frames][nlpf]. nlpf == 0 drives the
 * vocoder without LPF stream (excitation.rs:87-99; lpf may be NULL). An even nlpf > 0 is the
 * reference's panic (speech.rs:38-40) and returns JB_ERR_ARGUMENT. */
typedef struct jb_vocoder_config {
  size_t nmcp, nlpf, stage;
  int use_log_gain;
  size_t rate;
  double alpha, beta, volume;
  size_t fperiod;
  int coefficient_rows; /* test seam, normally 0. 1: `spectrum` holds filter coefficient rows instead,
                           [frames][nmcp] followed by the [n_utt][nmcp] rows the first frames start from
                           (the `coefficients` of vocoder/mod.rs:79-109); the spectrum transform is skipped */
} jb_vocoder_config;
/* One call: host tracks in, waveform out (library-owned buffer, jb_free; out_off[n_utt + 1] optional). */
int jb_engine_vocoder_synthesize(const jb_engine* e, const jb_vocoder_config* cfg, size_t n_utt,
                                 const size_t* frame_off, const double* spectrum, const double* lf0,
                                 const double* lpf, double** out, size_t* out_off);
/* The same as a resident batch for jb_resident_run / _fetch / _free (run() skips MLPG). */
int jb_resident_prepare_tracks(const jb_engine* e, const jb_vocoder_config* cfg, size_t n_utt, const size_t* frame_off,
                               const double* spectrum, const double* lf0, const double* lpf, int device_index,
                               jb_resident** out);
/* After jb_resident_run: the per-frame filter coefficient rows the run used, [frames][nmcp], then the
 * [n_utt][nmcp] rows the first frames started from (cap = doubles available in out). */
int jb_resident_rows(const jb_engine* e, jb_resident* r, double* out, size_t cap);
/* Runs MLPG + vocoder kernels; device time in ms by CUDA events; timings as above (optional). */
int jb_resident_run(const jb_engine* e, jb_resident* r, double* device_ms, double* timings);
int jb_resident_fetch(const jb_engine* e, jb_resident* r, double** out, size_t* out_off);
void jb_resident_free(const jb_engine* e, jb_resident* r);
/* Device-resident delivery (the launcher seam a GPU-side consumer binds; no 8 B/sample D2H copy).
 * jb_resident_device_output: the waveform of the last jb_resident_run as a DEVICE pointer, out_off[n_utt]
 * doubles on CUDA device *device, valid until the next run / free of `r`; the producing stream is
 * synchronised before the call returns, so any stream of the caller may read it.
 * jb_engine_synthesize_batch_device: Engine::synthesize for a batch (engine.rs:294-296) whose Vec<f64> stays
 * in HBM: prepare + run + device output in one call; release *keep with jb_resident_free when done. */
int jb_resident_device_output(const jb_engine* e, jb_resident* r, const double** d_out, size_t* out_off, int* device);
int jb_engine_synthesize_batch_device(const jb_engine* e, const char* const* lines, const size_t* line_off, size_t n_utt,
                                      int device_index, jb_resident** keep, const double** d_out, size_t* out_off,
                                      double* timings);
/* The MLPG seam below the engine: MlpgAdjust::new(gv_weight, msd_threshold, stream, gv, windows) + create(durations)
 * (src/mlpg_adjust/mod.rs:31-95) for ONE stream of n_utt utterances. state_off[n_utt + 1] (states, ascending from
 * 0); durations[states] (frames per state); windows as in Windows (src/model/voice/window.rs:4-56): n_windows odd
 * widths (<= 5, the static window first) and their coefficients concatenated; mean / vari [states][n_windows * dim]
 * with index dim * w + d (src/model/mod.rs:62); msd [states] or NULL (not an MSD stream: every frame is kept);
 * gv_mean / gv_vari [n_utt][dim] + gv_switch [states], or all NULL (no GV). *out: [frames][dim] row-major, -1e10
 * (NODATA) in the masked frames of an MSD stream; release with jb_free. frame_off[n_utt + 1] optional. */
typedef struct jb_mlpg_input {
  size_t n_utt, dim;
  const size_t* state_off;
  const int32_t* durations;
  size_t n_windows;
  const size_t* window_widths;
  const double* window_coefficients;
  const double *mean, *vari, *msd;
  double msd_threshold;
  const double *gv_mean, *gv_vari;
  const uint8_t* gv_switch;
  double gv_weight;
} jb_mlpg_input;
int jb_engine_mlpg_create(const jb_engine* e, const jb_mlpg_input* in, int device_index, double** out, size_t* frame_off);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The MLPG solve of jb_engine_mlpg_create (calc_wuw_and_wum + ldl_factorization + substitutions, same operation
 * order, same results) for ONE utterance of a stream that is not MSD and has no GV, with NO allocation: the
 * work planes live in a fixed 6 MiB `static __device__` array that every GPU loading this library holds until exit.
 * The call makes no cudaMalloc / cudaFree / allocating library call and needs no engine.
 *   device            CUDA device ordinal (not an index into an engine's device list); the caller's current
 *                     device is restored before returning
 *   d_mean, d_vari    DEVICE pointers, [n_frames][n_windows * dim], index dim * w + d: the state-level rows of
 *                     jb_mlpg_input repeated by their durations
 *   window_widths / window_coefficients   HOST pointers, as in jb_mlpg_input (n_windows <= 4, odd widths <= 5)
 *   d_out             DEVICE pointer, [n_frames][dim]
 *   stream            the caller's cudaStream_t (NULL: the default stream). The two kernels are only ENQUEUED on
 *                     it; the call does not synchronise.
 * Limit: n_frames * dim <= JB_MLPG_NOALLOC_MAX_CELLS, else JB_ERR_ARGUMENT (jb_last_error names the limit).
 * The array is shared by all calls on a device and keeps the last call's contents (every cell a call reads it
 * has written itself first, nothing needs clearing): calls whose kernels can overlap on one GPU - different
 * streams, or several host threads - race on it and must be serialised by the caller. */
#define JB_MLPG_NOALLOC_MAX_CELLS 131072
int jb_mlpg_solve_noalloc(int device, const double* d_mean, const double* d_vari, size_t n_frames, size_t dim,
                          size_t n_windows, const size_t* window_widths, const double* window_coefficients,
                          double* d_out, void* stream);

/* Process-wide knobs (no reference equivalent). jb_set_host_threads: host threads the front-end and the
 * batching layer use (0 = all cores; env JB_HOST_THREADS) - set it to cores / ranks when several processes
 * share a box. jb_trim_pools: frees the cached pinned host buffers (up to 24 GB are kept for reuse);
 * returns the bytes released. */
int jb_set_host_threads(int n);
/* Test hook (host only, needs no GPU): the time segments the planner would cut for utterances of the given frame
 * counts with `warm` frames of warm-up, `target` segments per filter launch and - when `gain` (one log-gain value per
 * frame, utterances concatenated) is given - the warm-up by loudness (jb_engine_set_mlsa_loud_warm). out: rows of
 * {utt, begin, end, warm, prev}; *n_segments rows exist (call with out = NULL to size the buffer). */
int jb_debug_plan_segments(const int64_t* frames, size_t n_utt, int warm, int warm_loud, double threshold,
                           const double* gain, int64_t target, int32_t* out, size_t cap, size_t* n_segments);
size_t jb_trim_pools(void);

/* Engine::generator + SpeechGenerator (src/engine.rs:301, src/speech.rs:25-96), batched:
 * MLPG runs at open; every step emits up to n_frames frames per utterance. */
/* `device`: index into the engine's device list (jb_engine_set_devices), as for jb_resident_prepare. */
int jb_stream_open(const jb_engine* e, const char* const* lines, const size_t* line_off,
                   size_t n_utt, int device, jb_stream** out);
size_t jb_stream_fperiod(const jb_stream* s);            /* SpeechGenerator::fperiod */
size_t jb_stream_synthesized_frames(const jb_stream* s); /* SpeechGenerator::synthesized_frames */
/* SpeechGenerator::generate_step for n_frames frames of every utterance. buf holds
 * n_utt * n_frames * fperiod doubles (cap = its length; too short -> JB_ERR_ARGUMENT, the
 * reference panics). written[u] = samples produced for utterance u, 0 once it has ended. */
int jb_stream_step(jb_stream* s, size_t n_frames, double* buf, size_t cap, size_t* written,
                   double* device_ms);
/* Page-locked host memory from the library's pool (no reference equivalent): a `buf` for jb_stream_step that the
 * GPU can write directly. With a pageable buffer every step's waveform is staged through the driver (4096 streams,
 * one frame: 0.71 ms per step instead of 0.47). Release with jb_free. */
int jb_alloc_pinned(size_t bytes, void** out);
void jb_stream_close(jb_stream* s);

#ifdef __cplusplus
}
#endif
#endif /* JBONSAI_B200_H_ */
