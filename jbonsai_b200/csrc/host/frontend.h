// Host front-end: labels, voice set, interpolation weights, Condition, state durations and the
// per-utterance parameter pack that is the *input contract of the hot path* (SURVEY.md §8a A0).
//
// Mirrors `src/label.rs`, `src/model/{mod,voice_set,interporation_weight,stream_parameter}.rs`,
// `src/duration.rs` and `Condition` in `src/engine.rs:32-243` of the reference. Everything here
// is host code by design (north star: label parsing, tree lookup, durations and the batching
// layer stay on the host).
#pragma once

#include <cfloat>
#include <string>
#include <utility>
#include <vector>

#include "voice.h"

namespace jb {

// src/constants.rs:4-13
constexpr double kMaxLf0 = 9.903487552536127;
constexpr double kMinLf0 = 2.995732273553991;
constexpr double kHalfTone = 0.05776226504666211;
constexpr double kDb = 0.11512925464970228;
constexpr double kNoData = -1e10;

// ---- labels (src/label.rs)
struct Labels {
  std::vector<std::string> labels;
  std::vector<std::pair<double, double>> times;  // frames; -1 = unspecified
  static Labels load_from_strings(size_t sampling_rate, size_t fperiod,
                                  const std::vector<std::string>& lines);
};

// ---- interpolation weights (src/model/interporation_weight.rs)
struct InterporationWeight {
  size_t nvoices = 1;
  std::vector<double> duration;
  std::vector<std::vector<double>> parameter;
  std::vector<std::vector<double>> gv;
  InterporationWeight() : InterporationWeight(1, 0) {}
  InterporationWeight(size_t nvoices, size_t nstream);
  void set_duration(const std::vector<double>& w);
  void set_parameter(size_t stream, const std::vector<double>& w);
  void set_gv(size_t stream, const std::vector<double>& w);

 private:
  void check(const std::vector<double>& w) const;  // throws kErrWeight
};

// ---- voice set (src/model/voice_set.rs)
struct VoiceSet {
  std::vector<std::shared_ptr<Voice>> voices;
  explicit VoiceSet(std::vector<std::shared_ptr<Voice>> v);  // throws kErrEmptyVoice/kErrMetadata
  const GlobalModelMetadata& global_metadata() const { return voices[0]->metadata; }
  const StreamModelMetadata& stream_metadata(size_t s) const {
    return voices[0]->stream_models[s].metadata;
  }
  const Windows& stream_windows(size_t s) const { return voices[0]->stream_models[s].windows; }
  size_t size() const { return voices.size(); }
};

// ---- Condition (src/engine.rs:32-243); setters clamp exactly like the reference.
struct Condition {
  size_t sampling_frequency = 0;
  size_t fperiod = 0;
  double volume = 1.0;
  std::vector<double> msd_threshold;
  std::vector<double> gv_weight;
  bool phoneme_alignment_flag = false;
  double speed = 1.0;
  size_t stage = 0;
  bool use_log_gain = false;
  double alpha = 0.0;
  double beta = 0.0;
  double additional_half_tone = 0.0;
  InterporationWeight interporation_weight;

  void load_model(const VoiceSet& voices);  // engine.rs:84-125
  void set_sampling_frequency(size_t i) { sampling_frequency = i < 1 ? 1 : i; }
  void set_fperiod(size_t i) { fperiod = i < 1 ? 1 : i; }
  void set_volume(double db);
  double get_volume() const;
  void set_msd_threshold(size_t s, double f);
  void set_gv_weight(size_t s, double f);
  void set_speed(double f) { speed = f > 1.0e-06 ? f : 1.0e-06; }
  void set_alpha(double f) { alpha = f < 0.0 ? 0.0 : (f > 1.0 ? 1.0 : f); }
  void set_beta(double f) { beta = f < 0.0 ? 0.0 : (f > 1.0 ? 1.0 : f); }
};

// Keyed access used by the C ABI ("speed", "volume", "alpha", "beta", "additional_half_tone",
// "msd_threshold"[idx], "gv_weight"[idx], "phoneme_alignment", "sampling_frequency", "fperiod",
// "stage", "use_log_gain"). Setters apply the reference's clamps. Throws kErrArgument.
void condition_set(Condition& c, const std::string& key, size_t idx, double value);
double condition_get(const Condition& c, const std::string& key, size_t idx);

// ---- durations (src/duration.rs)
std::vector<size_t> estimate_duration(const MeanVari* p, size_t n, double rho);
std::vector<size_t> estimate_duration_with_frame_length(const MeanVari* p, size_t n,
                                                        double frame_length);
std::vector<size_t> durations_create(const std::vector<MeanVari>& params, double speed);
std::vector<size_t> durations_create_with_alignment(
    const std::vector<MeanVari>& params, size_t nstate,
    const std::vector<std::pair<double, double>>& times);

// ---- Models (src/model/mod.rs:60-157): interpolated parameters for a label sequence.
struct Models {
  const std::vector<std::string>& labels;
  const VoiceSet& voices;
  const InterporationWeight& weights;
  size_t nstate() const { return voices.global_metadata().num_states; }
  std::vector<MeanVari> duration() const;
  // Per state: nwin*dim MeanVari (index dim*w + d) and msd (DBL_MAX when the stream has none).
  void stream(size_t s, std::vector<std::vector<MeanVari>>* params, std::vector<double>* msd) const;
  // Returns false when the stream has no GV (or there are no labels).
  bool gv(size_t s, std::vector<MeanVari>* params, std::vector<uint8_t>* gv_switch) const;
};

// ---- leaf pdfs of one stream, flattened for the device (the tables stay resident
// in HBM and the host sends one row index per (state, voice) instead of the interpolated vectors).
// Row r of voice v, tree t (state t + 2), pdf index i (1-based): voice_base[v] + tree_base[v][t] + i - 1.
struct StreamTable {
  size_t width = 0;   // nwin * dim
  bool is_msd = false;
  // false when some pdf disagrees with the stream metadata (length or MSD flag): the engine then
  // keeps interpolating on the host (VoiceSet::weighted) so that results stay exactly the reference's
  bool regular = true;
  std::vector<double> mean, vari;  // [rows][width]
  std::vector<double> msd;         // [rows] (is_msd only)
  std::vector<size_t> voice_base;
  std::vector<std::vector<size_t>> tree_base;
  size_t rows = 0;
};
std::vector<StreamTable> build_stream_tables(const VoiceSet& voices);

// ---- the hot path's input for one utterance (state level, before duration expansion)
struct StreamPack {
  size_t dim = 0;   // vector_length
  size_t nwin = 0;  // windows
  bool is_msd = false;
  std::vector<double> mean;  // [S][nwin*dim]
  std::vector<double> vari;  // [S][nwin*dim]
  std::vector<double> msd;   // [S] (DBL_MAX if !is_msd)
  // index mode (device-side gather): rows[i * nvoices + v] = StreamTable row of state i in voice v;
  // mean/vari/msd stay empty
  std::vector<int32_t> rows;
  bool has_gv = false;
  std::vector<double> gv_mean;     // [dim]  (already multiplied by nothing; gv_weight applied later)
  std::vector<double> gv_vari;     // [dim]
  std::vector<uint8_t> gv_switch;  // [S]
};

struct UtterancePack {
  size_t nstates = 0;  // S = labels * num_states
  size_t nframes = 0;  // T = sum(durations)
  std::vector<int32_t> durations;  // [S]
  std::vector<StreamPack> streams;
};

// Engine::generator up to (not including) MlpgAdjust::create  (engine.rs:301-357).
// tables != nullptr selects index mode for the streams whose table is regular.
UtterancePack build_utterance(const VoiceSet& voices, const Condition& cond,
                              const std::vector<std::string>& lines,
                              const std::vector<StreamTable>* tables = nullptr);

// Frames of one utterance: the label and duration part of build_utterance only (one tree walk per
// label instead of sixteen). Lets a batch lay out its output before the stream parameters exist.
size_t utterance_frames(const VoiceSet& voices, const Condition& cond, const std::vector<std::string>& lines);

}  // namespace jb
