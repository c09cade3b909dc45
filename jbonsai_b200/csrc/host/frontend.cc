// Host front-end implementation. Compile with -ffp-contract=off: the reference never fuses
// multiply-adds outside `mlsa::fir`, and interpolation / duration arithmetic must round alike.
#include "frontend.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace jb {

// ---------------------------------------------------------------- labels
namespace {

bool parse_f64(const std::string& s, double* out) {
  if (s.empty()) return false;
  char* e = nullptr;
  double v = strtod(s.c_str(), &e);
  if (e != s.c_str() + s.size()) return false;
  *out = v;
  return true;
}

// Structural check standing in for jlabel's parser: the HTS_TTS_JPN field separators must
// appear in order (p1^p2-p3+p4=p5/A:../B:.. … /K:..).
bool label_well_formed(const std::string& l) {
  static const char* seps[] = {"^", "-", "+", "=", "/A:", "/B:", "/C:", "/D:",
                               "/E:", "/F:", "/G:", "/H:", "/I:", "/J:", "/K:"};
  size_t pos = 0;
  for (const char* s : seps) {
    size_t p = l.find(s, pos);
    if (p == std::string::npos) return false;
    pos = p + strlen(s);
  }
  return true;
}

}  // namespace

Labels Labels::load_from_strings(size_t sampling_rate, size_t fperiod,
                                 const std::vector<std::string>& lines) {
  Labels out;
  // start/end times are in units of 100 ns (label.rs:46-47)
  const double rate = (double)sampling_rate / ((double)fperiod * 1e+7);
  for (const std::string& line : lines) {
    size_t sp1 = line.find(' ');
    if (sp1 == std::string::npos) {
      if (line.empty()) continue;
      if (!label_well_formed(line))
        throw Error(kErrLabelParse, "jlabel failed to parse fullcontext-label: " + line);
      out.times.push_back({-1.0, -1.0});
      out.labels.push_back(line);
      continue;
    }
    size_t sp2 = line.find(' ', sp1 + 1);
    if (sp2 == std::string::npos)
      throw Error(kErrLabelMissing, "Expected a fullcontext-label in " + line);
    double start, end;
    if (!parse_f64(line.substr(0, sp1), &start) ||
        !parse_f64(line.substr(sp1 + 1, sp2 - sp1 - 1), &end))
      throw Error(kErrLabelFloat, "Failed to parse as floating-point number");
    start *= rate;
    end *= rate;
    std::string label = line.substr(sp2 + 1);
    if (!label_well_formed(label))
      throw Error(kErrLabelParse, "jlabel failed to parse fullcontext-label: " + label);
    out.times.push_back({start, end});
    out.labels.push_back(label);
  }
  // Labels::new (label.rs:82-115)
  auto& t = out.times;
  for (size_t i = 0; i < t.size(); ++i) {
    if (i + 1 < t.size()) {
      if (t[i].second < 0.0 && t[i + 1].first >= 0.0)
        t[i].second = t[i + 1].first;
      else if (t[i].second >= 0.0 && t[i + 1].first < 0.0)
        t[i + 1].first = t[i].second;
    }
    if (t[i].first < 0.0) t[i].first = -1.0;
    if (t[i].second < 0.0) t[i].second = -1.0;
  }
  return out;
}

// ---------------------------------------------------------------- weights
InterporationWeight::InterporationWeight(size_t nv, size_t nstream) : nvoices(nv) {
  std::vector<double> avg(nv, 1.0 / (double)nv);
  duration = avg;
  parameter.assign(nstream, avg);
  gv.assign(nstream, avg);
}

void InterporationWeight::check(const std::vector<double>& w) const {
  double sum = 0.0;
  for (double x : w) sum += x;
  if (std::fabs(sum - 1.0) > DBL_EPSILON)  // approx::abs_diff_ne! default epsilon
    throw Error(kErrWeight, "Weights do not sum to 1.0");
  if (w.size() != nvoices) throw Error(kErrWeight, "Weights length is invalid");
}
void InterporationWeight::set_duration(const std::vector<double>& w) {
  check(w);
  duration = w;
}
void InterporationWeight::set_parameter(size_t s, const std::vector<double>& w) {
  check(w);
  if (s >= parameter.size()) throw Error(kErrArgument, "stream index out of range");
  parameter[s] = w;
}
void InterporationWeight::set_gv(size_t s, const std::vector<double>& w) {
  check(w);
  if (s >= gv.size()) throw Error(kErrArgument, "stream index out of range");
  gv[s] = w;
}

// ---------------------------------------------------------------- voice set
VoiceSet::VoiceSet(std::vector<std::shared_ptr<Voice>> v) : voices(std::move(v)) {
  if (voices.empty()) throw Error(kErrEmptyVoice, "No HTS voice was given.");
  const Voice& first = *voices[0];
  for (size_t i = 1; i < voices.size(); ++i) {
    const Voice& o = *voices[i];
    bool ok = o.metadata == first.metadata && o.stream_models.size() == first.stream_models.size();
    for (size_t s = 0; ok && s < o.stream_models.size(); ++s)
      ok = o.stream_models[s].metadata == first.stream_models[s].metadata;
    if (!ok) throw Error(kErrMetadata, "The global metadata does not match.");
  }
}

// ---------------------------------------------------------------- condition
void Condition::load_model(const VoiceSet& voices) {
  const auto& md = voices.global_metadata();
  size_t nstream = md.num_streams;
  sampling_frequency = md.sampling_frequency;
  fperiod = md.frame_period;
  msd_threshold.assign(nstream, 0.5);
  gv_weight.assign(nstream, 1.0);
  for (const std::string& option : voices.stream_metadata(0).option) {
    size_t eq = option.find('=');
    if (eq == std::string::npos) continue;  // "Skipped unrecognized option"
    std::string key = option.substr(0, eq), value = option.substr(eq + 1);
    if (key == "GAMMA") {
      bool ok = !value.empty();
      for (char c : value) ok = ok && c >= '0' && c <= '9';
      if (!ok) throw Error(kErrOption, "Failed to parse option GAMMA");
      stage = (size_t)strtoull(value.c_str(), nullptr, 10);
    } else if (key == "LN_GAIN") {
      if (value == "1")
        use_log_gain = true;
      else if (value == "0")
        use_log_gain = false;
      else
        throw Error(kErrOption, "Failed to parse option LN_GAIN");
    } else if (key == "ALPHA") {
      double a;
      if (!parse_f64(value, &a)) throw Error(kErrOption, "Failed to parse option ALPHA");
      alpha = a;
    }
  }
  interporation_weight = InterporationWeight(voices.size(), nstream);
}

void Condition::set_volume(double db) { volume = std::exp(db * kDb); }
double Condition::get_volume() const { return std::log(volume) / kDb; }
void Condition::set_msd_threshold(size_t s, double f) {
  if (s >= msd_threshold.size()) throw Error(kErrArgument, "stream index out of range");
  msd_threshold[s] = f < 0.0 ? 0.0 : (f > 1.0 ? 1.0 : f);
}
void Condition::set_gv_weight(size_t s, double f) {
  if (s >= gv_weight.size()) throw Error(kErrArgument, "stream index out of range");
  gv_weight[s] = f > 0.0 ? f : 0.0;
}

void condition_set(Condition& c, const std::string& key, size_t idx, double v) {
  if (key == "speed") c.set_speed(v);
  else if (key == "volume") c.set_volume(v);
  else if (key == "alpha") c.set_alpha(v);
  else if (key == "beta") c.set_beta(v);
  else if (key == "additional_half_tone") c.additional_half_tone = v;
  else if (key == "msd_threshold") c.set_msd_threshold(idx, v);
  else if (key == "gv_weight") c.set_gv_weight(idx, v);
  else if (key == "phoneme_alignment") c.phoneme_alignment_flag = v != 0.0;
  else if (key == "sampling_frequency") c.set_sampling_frequency(v < 0 ? 0 : (size_t)v);
  else if (key == "fperiod") c.set_fperiod(v < 0 ? 0 : (size_t)v);
  else if (key == "stage") c.stage = v < 0 ? 0 : (size_t)v;
  else if (key == "use_log_gain") c.use_log_gain = v != 0.0;
  else throw Error(kErrArgument, "unknown condition key " + key);
}

double condition_get(const Condition& c, const std::string& key, size_t idx) {
  if (key == "speed") return c.speed;
  if (key == "volume") return c.get_volume();
  if (key == "alpha") return c.alpha;
  if (key == "beta") return c.beta;
  if (key == "additional_half_tone") return c.additional_half_tone;
  if (key == "msd_threshold" && idx < c.msd_threshold.size()) return c.msd_threshold[idx];
  if (key == "gv_weight" && idx < c.gv_weight.size()) return c.gv_weight[idx];
  if (key == "phoneme_alignment") return c.phoneme_alignment_flag ? 1.0 : 0.0;
  if (key == "sampling_frequency") return (double)c.sampling_frequency;
  if (key == "fperiod") return (double)c.fperiod;
  if (key == "stage") return (double)c.stage;
  if (key == "use_log_gain") return c.use_log_gain ? 1.0 : 0.0;
  throw Error(kErrArgument, "unknown condition key " + key);
}

// ---------------------------------------------------------------- durations
static inline size_t round_to_usize_min1(double x) {
  double r = std::fmax(std::round(x), 1.0);  // NaN -> 1.0 like f64::max
  if (r >= 1.8446744073709552e19) return SIZE_MAX;
  return (size_t)r;
}

std::vector<size_t> estimate_duration(const MeanVari* p, size_t n, double rho) {
  std::vector<size_t> d(n);
  for (size_t i = 0; i < n; ++i) d[i] = round_to_usize_min1(p[i].mean + rho * p[i].vari);
  return d;
}

static inline uint64_t total_order_key_nonneg(double x) {
  uint64_t u;
  memcpy(&u, &x, 8);
  return u;  // x = |..| so sign bit is clear; bit pattern is monotone incl. +inf < +NaN
}

std::vector<size_t> estimate_duration_with_frame_length(const MeanVari* p, size_t n,
                                                        double frame_length) {
  size_t target = round_to_usize_min1(frame_length);
  if (target <= n) return std::vector<size_t>(n, 1);
  double mean = 0.0, vari = 0.0;
  for (size_t i = 0; i < n; ++i) {
    mean = mean + p[i].mean;
    vari = vari + p[i].vari;
  }
  double rho = ((double)target - mean) / vari;
  std::vector<size_t> d = estimate_duration(p, n, rho);
  if (d.empty()) return d;
  size_t sum = 0;
  for (size_t x : d) sum += x;
  auto cost = [&](size_t dur, const MeanVari& mv) {
    return std::fabs(rho - ((double)dur - mv.mean) / mv.vari);
  };
  while (target != sum) {
    if (target > sum) {
      size_t best = 0;
      uint64_t bk = total_order_key_nonneg(cost(d[0] + 1, p[0]));
      for (size_t i = 1; i < n; ++i) {
        uint64_t k = total_order_key_nonneg(cost(d[i] + 1, p[i]));
        if (k < bk) {
          bk = k;
          best = i;
        }
      }
      d[best] += 1;
      sum += 1;
    } else {
      long best = -1;
      uint64_t bk = 0;
      for (size_t i = 0; i < n; ++i) {
        if (d[i] <= 1) continue;
        uint64_t k = total_order_key_nonneg(cost(d[i] - 1, p[i]));
        if (best < 0 || k < bk) {
          bk = k;
          best = (long)i;
        }
      }
      if (best < 0) throw Error(kErrArgument, "duration fitting found no adjustable state");
      d[(size_t)best] -= 1;
      sum -= 1;
    }
  }
  return d;
}

std::vector<size_t> durations_create(const std::vector<MeanVari>& params, double speed) {
  std::vector<size_t> d = estimate_duration(params.data(), params.size(), 0.0);
  if (speed != 1.0) {
    size_t length = 0;
    for (size_t x : d) length += x;
    d = estimate_duration_with_frame_length(params.data(), params.size(), (double)length / speed);
  }
  return d;
}

std::vector<size_t> durations_create_with_alignment(
    const std::vector<MeanVari>& params, size_t nstate,
    const std::vector<std::pair<double, double>>& times) {
  std::vector<size_t> duration;
  size_t frame_count = 0, next_state = 0, state = 0;
  for (size_t i = 0; i < times.size(); ++i) {
    double end_frame = times[i].second;
    if (end_frame >= 0.0) {
      std::vector<size_t> cur = estimate_duration_with_frame_length(
          params.data() + next_state, state + nstate - next_state, end_frame - (double)frame_count);
      for (size_t x : cur) frame_count += x;
      next_state = state + nstate;
      duration.insert(duration.end(), cur.begin(), cur.end());
    }
    // (final label without an end time: the reference only prints a warning, duration.rs:57-60)
    state += nstate;
  }
  return duration;
}

// ---------------------------------------------------------------- models
namespace {

// VoiceSet::weighted (voice_set.rs:80-96): first*w0, then += w*rhs (separate mul and add).
template <class F>
ModelParameter weighted(const VoiceSet& vs, const std::vector<double>& w, F get) {
  const ModelParameter& first = get(*vs.voices[0]);
  ModelParameter r;
  r.parameters.resize(first.parameters.size());
  for (size_t i = 0; i < first.parameters.size(); ++i)
    r.parameters[i] = {first.parameters[i].mean * w[0], first.parameters[i].vari * w[0]};
  r.has_msd = first.has_msd;
  if (first.has_msd) r.msd = w[0] * first.msd;
  size_t n = vs.voices.size() < w.size() ? vs.voices.size() : w.size();
  for (size_t v = 1; v < n; ++v) {
    const ModelParameter& rhs = get(*vs.voices[v]);
    size_t m = r.parameters.size() < rhs.parameters.size() ? r.parameters.size() : rhs.parameters.size();
    for (size_t i = 0; i < m; ++i) {
      r.parameters[i].mean += w[v] * rhs.parameters[i].mean;
      r.parameters[i].vari += w[v] * rhs.parameters[i].vari;
    }
    if (r.has_msd && rhs.has_msd) r.msd += w[v] * rhs.msd;
  }
  return r;
}

}  // namespace

std::vector<MeanVari> Models::duration() const {
  std::vector<MeanVari> out;
  for (const std::string& label : labels) {
    ModelParameter p = weighted(voices, weights.duration, [&](const Voice& v) -> const ModelParameter& {
      return v.duration_model.get_parameter(2, label);
    });
    out.insert(out.end(), p.parameters.begin(), p.parameters.end());
  }
  return out;
}

void Models::stream(size_t s, std::vector<std::vector<MeanVari>>* params,
                    std::vector<double>* msd) const {
  size_t ns = nstate();
  params->clear();
  msd->clear();
  for (const std::string& label : labels)
    for (size_t st = 2; st < 2 + ns; ++st) {
      ModelParameter p =
          weighted(voices, weights.parameter[s], [&](const Voice& v) -> const ModelParameter& {
            return v.stream_models[s].stream_model.get_parameter(st, label);
          });
      msd->push_back(p.has_msd ? p.msd : DBL_MAX);
      params->push_back(std::move(p.parameters));
    }
}

bool Models::gv(size_t s, std::vector<MeanVari>* params, std::vector<uint8_t>* gv_switch) const {
  if (!voices.stream_metadata(s).use_gv) return false;
  if (labels.empty()) return false;
  const std::string& label = labels[0];
  ModelParameter p = weighted(voices, weights.gv[s], [&](const Voice& v) -> const ModelParameter& {
    if (!v.stream_models[s].has_gv) throw Error(kErrIndex, "voice has no GV model");
    return v.stream_models[s].gv_model.get_parameter(2, label);
  });
  *params = p.parameters;
  gv_switch->clear();
  size_t ns = nstate();
  for (const std::string& l : labels) {
    uint8_t sw = !voices.global_metadata().gv_off_context.test(l);
    gv_switch->insert(gv_switch->end(), ns, sw);
  }
  return true;
}

// ---------------------------------------------------------------- utterance pack
std::vector<StreamTable> build_stream_tables(const VoiceSet& voices) {
  const size_t nstream = voices.global_metadata().num_streams;
  std::vector<StreamTable> out(nstream);
  for (size_t s = 0; s < nstream; ++s) {
    const StreamModelMetadata& md = voices.stream_metadata(s);
    StreamTable& t = out[s];
    t.width = md.vector_length * voices.stream_windows(s).size();
    t.is_msd = md.is_msd;
    for (const auto& v : voices.voices) {
      const Model& m = v->stream_models[s].stream_model;
      t.voice_base.push_back(t.rows);
      std::vector<size_t> tb;
      size_t r = 0;
      for (const auto& tree_pdfs : m.pdf) {
        tb.push_back(r);
        r += tree_pdfs.size();
      }
      t.tree_base.push_back(std::move(tb));
      t.rows += r;
    }
    t.mean.resize(t.rows * t.width);
    t.vari.resize(t.rows * t.width);
    if (t.is_msd) t.msd.resize(t.rows);
    size_t row = 0;
    for (const auto& v : voices.voices)
      for (const auto& tree_pdfs : v->stream_models[s].stream_model.pdf)
        for (const ModelParameter& p : tree_pdfs) {
          if (p.parameters.size() != t.width || p.has_msd != t.is_msd) {
            t.regular = false;
          } else {
            for (size_t k = 0; k < t.width; ++k) {
              t.mean[row * t.width + k] = p.parameters[k].mean;
              t.vari[row * t.width + k] = p.parameters[k].vari;
            }
            if (t.is_msd) t.msd[row] = p.msd;
          }
          ++row;
        }
    if (t.rows >= (size_t)INT32_MAX) t.regular = false;
  }
  return out;
}

size_t utterance_frames(const VoiceSet& voices, const Condition& cond, const std::vector<std::string>& lines) {
  Labels labels = Labels::load_from_strings(cond.sampling_frequency, cond.fperiod, lines);
  Models models{labels.labels, voices, cond.interporation_weight};
  std::vector<MeanVari> dur = models.duration();
  std::vector<size_t> durations = cond.phoneme_alignment_flag
                                      ? durations_create_with_alignment(dur, models.nstate(), labels.times)
                                      : durations_create(dur, cond.speed);
  size_t n = 0;
  for (size_t d : durations) n += d;
  return n;
}

UtterancePack build_utterance(const VoiceSet& voices, const Condition& cond,
                              const std::vector<std::string>& lines,
                              const std::vector<StreamTable>* tables) {
  Labels labels = Labels::load_from_strings(cond.sampling_frequency, cond.fperiod, lines);
  Models models{labels.labels, voices, cond.interporation_weight};
  std::vector<MeanVari> dur = models.duration();
  std::vector<size_t> durations = cond.phoneme_alignment_flag
                                      ? durations_create_with_alignment(dur, models.nstate(), labels.times)
                                      : durations_create(dur, cond.speed);
  UtterancePack pack;
  pack.nstates = durations.size();
  pack.durations.resize(durations.size());
  for (size_t i = 0; i < durations.size(); ++i) {
    pack.durations[i] = (int32_t)durations[i];
    pack.nframes += durations[i];
  }
  size_t nstream = voices.global_metadata().num_streams;
  std::vector<uint64_t> label_hashes;
  if (tables) {
    label_hashes.resize(labels.labels.size());
    for (size_t l = 0; l < labels.labels.size(); ++l) label_hashes[l] = label_hash(labels.labels[l]);
  }
  for (size_t s = 0; s < nstream; ++s) {
    const StreamModelMetadata& md = voices.stream_metadata(s);
    StreamPack sp;
    sp.dim = md.vector_length;
    sp.nwin = voices.stream_windows(s).size();
    sp.is_msd = md.is_msd;
    if (tables && (*tables)[s].regular) {
      // index mode: the same tree walks as Models::stream (model/mod.rs:98-118), rows instead of
      // vectors; the weighted sum of voice_set.rs:80-96 runs on the device (k_gather_states)
      const StreamTable& tb = (*tables)[s];
      const size_t nv = std::min(voices.size(), cond.interporation_weight.parameter[s].size());
      const size_t ns = models.nstate();
      const size_t S = std::min(pack.nstates, labels.labels.size() * ns);
      sp.rows.resize(S * nv);
      // one cache lookup per (label, voice) answers all `ns` states (Model::get_all), with the label hashed once
      // for all streams (label_hashes above)
      std::vector<int32_t> leaf;
      for (size_t v = 0; v < nv; ++v) {
        const Model& m = voices.voices[v]->stream_models[s].stream_model;
        std::vector<long> tpos(ns, -1);  // tree of state 2 + k (Model::get_index's search, voice/model.rs:50-67)
        for (size_t k = 0; k < ns; ++k)
          for (size_t t = 0; t < m.trees.size(); ++t)
            if (m.trees[t].state == 2 + k) {
              tpos[k] = (long)t;
              break;
            }
        leaf.assign(m.trees.size(), 0);
        for (size_t l = 0; l * ns < S; ++l) {
          const std::string& label = labels.labels[l];
          m.get_all(label, label_hashes[l], leaf.data());
          for (size_t k = 0; k < ns && l * ns + k < S; ++k) {
            const long ti = tpos[k];
            const long pi = ti < 0 ? -1 : (long)leaf[(size_t)ti];
            if (ti < 0 || pi < 1 || (size_t)ti >= m.pdf.size() || (size_t)(pi - 1) >= m.pdf[(size_t)ti].size())
              throw Error(kErrIndex, "tree/pdf index not found for label: " + label);
            sp.rows[(l * ns + k) * nv + v] = (int32_t)(tb.voice_base[v] + tb.tree_base[v][(size_t)ti] + (size_t)pi - 1);
          }
        }
      }
      std::vector<MeanVari> gvp;
      std::vector<uint8_t> gsw;
      if (models.gv(s, &gvp, &gsw)) {
        sp.has_gv = true;
        sp.gv_mean.resize(sp.dim);
        sp.gv_vari.resize(sp.dim);
        for (size_t d = 0; d < sp.dim; ++d) {
          sp.gv_mean[d] = gvp[d].mean;
          sp.gv_vari[d] = gvp[d].vari;
        }
        gsw.resize(S);
        sp.gv_switch = gsw;
      }
      pack.streams.push_back(std::move(sp));
      continue;
    }
    std::vector<std::vector<MeanVari>> params;
    std::vector<double> msd;
    models.stream(s, &params, &msd);
    // With phoneme alignment the reference zips streams with a possibly shorter duration
    // vector; keep only the states that have a duration.
    size_t S = pack.nstates < params.size() ? pack.nstates : params.size();
    size_t L = sp.dim * sp.nwin;
    sp.mean.resize(S * L);
    sp.vari.resize(S * L);
    sp.msd.assign(msd.begin(), msd.begin() + S);
    for (size_t i = 0; i < S; ++i) {
      if (params[i].size() < L) throw Error(kErrParse, "pdf shorter than vector_length*num_windows");
      for (size_t k = 0; k < L; ++k) {
        sp.mean[i * L + k] = params[i][k].mean;
        sp.vari[i * L + k] = params[i][k].vari;
      }
    }
    if (s == 1 && cond.additional_half_tone != 0.0) {
      // StreamParameter::apply_additional_half_tone (stream_parameter.rs:29-37)
      for (size_t i = 0; i < S; ++i) {
        double m = sp.mean[i * L] + cond.additional_half_tone * kHalfTone;
        sp.mean[i * L] = m < kMinLf0 ? kMinLf0 : (m > kMaxLf0 ? kMaxLf0 : m);
      }
    }
    std::vector<MeanVari> gvp;
    std::vector<uint8_t> gsw;
    if (models.gv(s, &gvp, &gsw)) {
      sp.has_gv = true;
      sp.gv_mean.resize(sp.dim);
      sp.gv_vari.resize(sp.dim);
      for (size_t d = 0; d < sp.dim; ++d) {
        sp.gv_mean[d] = gvp[d].mean;
        sp.gv_vari[d] = gvp[d].vari;
      }
      gsw.resize(S);
      sp.gv_switch = gsw;
    }
    pack.streams.push_back(std::move(sp));
  }
  return pack;
}

}  // namespace jb
