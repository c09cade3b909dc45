// Device-side data layout and kernel launchers for the hot path (sm_100a).
//
// Path: MlpgAdjust::create (src/mlpg_adjust/mod.rs:51-95) for the MCP / LF0 /
// LPF streams of every utterance of a batch, then Vocoder::synthesize (src/vocoder/mod.rs:72-178)
// over every frame. All launchers are asynchronous on `stream` and return cudaError_t as int.
#pragma once

#include <cstddef>
#include <cstdint>

#include <cuda_runtime.h>

namespace jbk {

constexpr int kMaxWin = 4;       // windows per stream (static, delta, delta-delta, +1)
constexpr int kMaxWinWidth = 5;  // taps per window
constexpr int kPade = 5;         // MLSA Pade order (N = 6, src/vocoder/mlsa.rs:31)

struct WindowSet {
  int nwin;
  int width[kMaxWin];
  double coef[kMaxWin][kMaxWinWidth];
  int band_width() const {  // MlpgMatrix width = max_width*2+1 (mlpg.rs:27)
    int m = 0;
    for (int i = 0; i < nwin; ++i) m = width[i] > m ? width[i] : m;
    return (m / 2) * 2 + 1;
  }
};

// One batch, all utterances concatenated. "states" = labels * num_states, "frames" = sum of
// state durations, "samples" = frames * fperiod.
struct BatchLayout {
  int n_utt;
  const int64_t* state_off;  // [n_utt+1]
  const int64_t* frame_off;  // [n_utt+1]
  const int32_t* durations;  // [total_states]
  int32_t* frame_state;      // [total_frames]  (out of expand) state index within the utterance
  int max_frames;            // longest utterance of the batch (frames)
  const int32_t* order;      // [n_utt] utterances by decreasing length (block scheduling order) or null
};

// Per stream state-level input (A0 contract) + MLPG work arrays.
struct StreamDev {
  int dim;
  int nwin;  // == windows.nwin
  int is_msd;
  int has_gv;
  WindowSet windows;
  double msd_threshold;
  double gv_weight;
  const double* mean;        // [total_states][nwin*dim]
  const double* vari;        // [total_states][nwin*dim]
  double* ivar;              // [total_states][nwin*dim] work: MeanVari::with_ivar(vari) (split path)
  const double* msd;         // [total_states] (only read when is_msd)
  const double* gv_mean;     // [n_utt][dim]
  const double* gv_vari;     // [n_utt][dim]
  const uint8_t* gv_switch;  // [total_states]
  // MSD bookkeeping (only when is_msd): all [total_frames] / [n_utt]
  int32_t* left;     // distance to the left mask edge   (mask.rs:51-82)
  int32_t* right;    // distance to the right mask edge
  int32_t* vidx;     // compacted index -> frame (within utterance)
  int32_t* nvoiced;  // [n_utt] number of unmasked frames
  // scratch planes, each [total_frames][dim]
  double* scratch;   // n_scratch_planes(band) planes
  double* par;       // output [total_frames][dim] (NODATA in masked frames)
};

inline int mlpg_scratch_planes(int band_width, int has_gv) {
  // factorised band (bw) + g + W'U^-1 mu + un-factorised band (bw) [+ solution + gradient for GV];
  // see the plane map in mlpg.cu (the general kernel uses a subset of this)
  return 2 * band_width + 2 + (has_gv ? 2 : 0);
}

struct StreamSet {
  int n;
  StreamDev s[3];
};

// Device-side parameter gather: the leaf pdfs of a stream stay resident in HBM
// (`tab_*`, [rows][width]); the host sends rows[state][nvoices] and the kernel forms the
// interpolated state-level vectors exactly like VoiceSet::weighted (src/model/voice_set.rs:80-96:
// first * w0, then += w_v * rhs with separate multiply and add), plus
// StreamParameter::apply_additional_half_tone (src/model/stream_parameter.rs:29-37) when asked.
constexpr int kMaxVoices = 8;
struct GatherStream {
  int width, is_msd, nvoices;
  double w[kMaxVoices];
  const double* tab_mean;
  const double* tab_vari;
  const double* tab_msd;
  const int32_t* rows;  // [total_states][nvoices]
  double* mean;         // [total_states][width]
  double* vari;
  double* msd;          // [total_states] (is_msd only)
  int half_tone;        // apply mean[0] = clamp(mean[0] + half_tone_add)
  double half_tone_add, lf0_min, lf0_max;
};
int launch_gather_states(const GatherStream& g, int64_t total_states, cudaStream_t stream);

int launch_expand(const BatchLayout& b, const StreamSet& streams, cudaStream_t stream);
int launch_mlpg(const BatchLayout& b, const StreamDev& s, int64_t total_frames, int64_t total_states,
                cudaStream_t stream);

struct VocoderParams {
  int nmcp;     // 35
  int nlpf;     // 31 (0 = no LPF stream)
  int fperiod;  // 240
  double rate;  // 48000
  double alpha;
  double volume;
  int stage = 0;         // 0: mel-cepstrum in, MLSA filter; n > 0: line spectral pairs in, MGLSA filter of n
                         // sections, gamma = -1/n (stage.rs:28-38)
  int use_log_gain = 0;  // stage > 0: the gain slot holds ln(gain) (lsp.rs:92-96)
};

// Per-utterance vocoder state carried between chunked launches (streaming, speech.rs:65-83).
// Layout: [n_utt][kVocStateDoubles].
//   [0..170)   d21[5][34]  (taps d[1..34] of each Pade stage; d[0] is write-only in the reference)
//   [170..176) d11, [176..182) d12, [182..188) d22   (mlsa.rs:5-10)
//   [188..223) coefficients[35]                      (vocoder/mod.rs Stage::Zero)
//   [223] pitch_of_curr_point [224] pitch_counter    (excitation.rs:2-10)
//   [226..251) variant 2 only: carry-in of tap block b = 1..5 of stage s at [226 + 5 s + b - 1]
//              (== d21[s][6b-1] up to rounding; kept so that chunked == whole-utterance bit for bit)
constexpr int kVocStateDoubles = 256;

struct VocoderDev {
  const double* mcp;    // [total_frames][nmcp] MLPG output
  const double* lf0;    // [total_frames]
  const double* lpf;    // [total_frames][nlpf]
  const double* noise;  // [>= max samples per utterance] global Gaussian sequence
  double* b;            // [total_frames][nmcp]  mc2b(mcp), post-filtered when beta > 0 (out of frame_prep / postfilter)
  const double* b_first;  // [n_utt][nmcp] or nullptr. beta > 0 only: mc2b of the un-filtered first frame, the
                          // coefficients the reference starts from (is_first, vocoder/mod.rs:79-88); nullptr =
                          // row 0 of `b` (identical when beta == 0)
  double* binc;         // [total_frames][nmcp] per-sample coefficient increments of every frame
                        // (launch_row_increments): read by the segmented filters (MLSA variant 3, MGLSA); else nullptr
  double* pitch;        // [total_frames]        p per frame (out of frame_prep)
  double* pulse;        // [total_samples]       pulse amplitude or 0 (out of pitch kernel)
  int32_t* runs;        // [1 + total_frames]    work: number of voiced runs, then their first frames
  double* wave;         // [total_samples]       excitation*gain; variants 0-2 filter it in place
  double* out;          // [total_samples]       variant 3 only: filtered waveform (wave stays the input)
  double* state;        // [n_utt][kVocStateDoubles] or nullptr (whole-utterance synthesis)
  int frame_begin;      // first frame (per utterance) processed by this launch
  int frame_count;      // frames processed per utterance by this launch (clipped to its length)
};

// what: bit 0 = pitch period from lf0, bit 1 = b = mc2b(mcp)
int launch_frame_prep(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d,
                      int64_t total_frames, int what, cudaStream_t stream);
// what = 2 and launch_row_increments (no b_first) in one pass; stage 0, rows of up to 47 coefficients.
bool coef_rows_supported(const VocoderParams& p);
int launch_coef_rows(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, int64_t total_frames,
                     cudaStream_t stream);
// beta > 0 (MelCepstrum::postfilter_mcp, cepstrum.rs:23-37): first copies row 0 of every utterance
// of d.b (= mc2b of the raw first frame) to b_first[n_utt][nmcp], then rewrites every row of d.b as
// mc2b(postfilter_mcp(c)). Runs after launch_frame_prep(what & 2). No-op when beta <= 0 or nmcp <= 2.
int launch_postfilter(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, double beta,
                      double* b_first, int64_t total_frames, cudaStream_t stream);
int launch_pulse(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, int64_t total_samples,
                 cudaStream_t stream);
int launch_excitation(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d,
                      int64_t total_frames, cudaStream_t stream);
// variant: 0 = reference operation order, stage-per-lane; anything else = software-pipelined blocked scan
// (mlsa.cu, one warp per utterance). Vector lengths other than 35, frame periods that are no multiple of 16 and
// alpha ~ 0 run the generic-length kernel (k_mlsa_any: any 2 <= nmcp <= 128, any fperiod).
int launch_mlsa_v2(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, cudaStream_t stream);
int launch_mlsa(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, int variant,
                cudaStream_t stream);

// Variant 3 (mlsa_seg.cu): utterances cut into time segments that are filtered independently.
// Segment = frames [begin, end) of utterance `utt`; filtering starts at frame `warm` <= begin from a
// zero state (warm == begin == 0 for the first segment) and the output of [warm, begin) is discarded.
// `prev` = id of the segment that ends where this one begins (-1 for the first of an utterance).
struct MlsaSegment {
  int32_t utt, begin, end, warm, prev;
};
constexpr int kMlsaSegStateDoubles = 192;  // per (segment, side): 5 stages x (34 taps, d11, d12, d22)
struct MlsaSegments {
  int n_seg;
  const MlsaSegment* seg;  // [n_seg], six consecutive segments share a warp (host sorts by length)
  double* state;           // [n_seg][2][kMlsaSegStateDoubles] cut-check taps, or nullptr
  double* mismatch;        // [n_seg] or nullptr: mismatch of every segment's cut (out of launch_mlsa_seg_check)
  // Repair pass (launch_mlsa_v3 only): a segment with warm == begin and prev >= 0 starts from the end state of
  // segment `prev` in this array (the state array of the original pass) instead of a warm-up from zero.
  const double* init_state = nullptr;
  // Cuts that failed the check, compacted on the device: ids into `seg`, count (may exceed fail_cap: then the
  // list is incomplete and the caller falls back), and a pinned host word the count is published to.
  int* fail_count = nullptr;
  int* fail_list = nullptr;
  int fail_cap = 0;
  int* host_fail_count = nullptr;
};
// Filters segments [seg_first, seg_first + seg_count) of d.wave into d.out (one launch = one wave).
// Needs d.binc (launch_row_increments), nmcp for which mlsa_v3_supports() is true, an even fperiod.
bool mlsa_v3_supports(int nmcp);
int mlsa_v3_chunk(int fperiod);  // samples per staged chunk; 0 = fperiod not supported (odd)
int launch_mlsa_v3(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, const MlsaSegments& sg,
                   int seg_first, int seg_count, cudaStream_t stream);
// After every wave has run: the largest state mismatch over all cuts -> *worst (device), for the
// host to compare with its tolerance.
// `host_mirror` (pinned, device-visible, or nullptr): also stored there by the device, no copy engine.
// `stride` / `count`: doubles per (segment, side) slot and how many of them are compared. `relative`:
// every cut's mismatch is divided by max(1, largest |state value| at that cut) (MGLSA: the output scale
// of an LSP voice is not known in advance). `prev_state`: the state array the `prev` ids of `sg` index when it is
// not sg.state (a repair pass over a few segments, compared with the end states of the original pass).
// `gain` (MLSA): the mismatch of a cut is scaled to a bound on the waveform error, mismatch * max(1, sum_k |c_k|
// of the coefficient rows at the cut) * |volume|, and compared with gain->tol on the device (sg.fail_*).
struct SegCheckGain {
  const BatchLayout* layout;
  const double* rows;  // VocoderDev::b
  int nm;
  double volume, tol;
};
int launch_mlsa_seg_check(const MlsaSegments& sg, double* worst, double* host_mirror, cudaStream_t stream,
                          int stride = kMlsaSegStateDoubles, int count = 5 * 37, bool relative = false,
                          const double* prev_state = nullptr, const SegCheckGain* gain = nullptr);
// Repair pass epilogue: segment i of fix[n_fix] (device) re-filtered frames [begin, end) from the exact state
// into `scratch` (indexed like `out`); where the last re-filtered frame agrees with `out` to `tol` the new
// samples replace the old ones and ok[i] = 1, else ok[i] = 0 and `out` stands.
int launch_mlsa_repair_merge(const BatchLayout& b, const MlsaSegment* fix, int n_fix, const double* scratch, double* out,
                             int fperiod, double tol, int* ok, cudaStream_t stream);

// Stage != 0 (mglsa.cu). launch_lsp_prep: rows of d.b = filter coefficients of every LSP frame of d.mcp
// (postfilter_lsp(beta), check_lsp_stability, lsp2mgc, mc2b, gnorm, [1..] *= gamma; vocoder/mod.rs:148-155)
// and b_first[n_utt][nmcp] = the same transform of every first frame without post-filter and stability
// check (mod.rs:92-107). nmcp <= 64.
int launch_lsp_prep(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, double beta, double* b_first,
                    int64_t total_frames, cudaStream_t stream);
// MGLSA filter (mglsa.rs:15-41) over time segments like launch_mlsa_v3: d.wave (excitation times
// coefficients[0]) -> d.out. 32 / stage segments share a warp. nmcp <= kMglsaMaxOrder, stage <= 8,
// fperiod % 16 == 0. Cut-check state: mglsa_state_doubles(stage) doubles per (segment, side).
// After the rows of d.b / d.b_first are final: d.binc[f][k] = (b[f][k] - row before)[k] / fperiod (vocoder/mod.rs:156-160).
int launch_row_increments(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, int64_t total_frames,
                          cudaStream_t stream);
constexpr int kMglsaMaxOrder = 36;
constexpr int kMglsaMaxStage = 8;
constexpr int mglsa_state_doubles(int stage) { return stage * kMglsaMaxOrder; }
// `literal`: the tap update as the reference writes it, d[i] += alpha * (d[i+1] - d[i-1]) with separate
// multiplies and adds (bit-identical to the oracle given the same rows, three dependent operations per tap);
// otherwise (d[i] + alpha d[i+1]) - alpha d'[i-1] with fused multiply-adds, one operation on the serial path.
int launch_mglsa(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, const MlsaSegments& sg,
                 int seg_first, int seg_count, bool literal, cudaStream_t stream);

// Sequential MGLSA with carried state (generate_step with stage != 0; also any frame period): filters frames
// [d.frame_begin, + d.frame_count) of every utterance in place in d.wave. `state`: [n_utt][kMglsaMaxStage][kMglsaMaxOrder]
// delay lines carried between launches, or nullptr (whole utterances from a zero state).
int launch_mglsa_stream(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d, double* state,
                        cudaStream_t stream);

// Streaming: copy frames [frame_begin, frame_begin+frame_count) of every utterance into
// chunk[n_utt][frame_count*fperiod] (zero padded past the end of an utterance).
int launch_gather_chunk(const BatchLayout& b, const VocoderParams& p, const VocoderDev& d,
                        double* chunk, cudaStream_t stream);

// f64 samples -> 16-bit PCM with the reference examples' convention (clamp to the i16 range, then
// truncate toward zero; examples/genji/main.rs:34-35). n must be a multiple of 4.
int launch_pcm16(const double* in, int16_t* out, int64_t n, cudaStream_t stream);

// mlpg_noalloc.cu: MLPG solve of one non-MSD stream with its work planes in a fixed __device__ array (no
// allocation of any kind). frames x dim of one call is bounded by kMlpgNoAllocMaxCells (JB_MLPG_NOALLOC_MAX_CELLS).
constexpr long long kMlpgNoAllocMaxCells = 131072;
int mlpg_solve_noalloc(int device, const double* d_mean, const double* d_vari, long long n_frames, int dim,
                       const WindowSet& ws, double* d_out, cudaStream_t stream, const char** what);

// Roofline denominators measured on the current device (peaks.cu); not on the synthesis path.
int measure_peaks(double* fp64_fma_tflops, double* hbm_copy_gbs);

}  // namespace jbk
