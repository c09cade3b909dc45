// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
// C entry points (ctypes) that run the whole reference pipeline on the CPU:
//   shared host front-end (jbonsai_b200/csrc/host, = the reference's out-of-scope host code)
//   -> oracle::mlpg_create -> oracle::Vocoder           (= Engine::synthesize, engine.rs:294-366)
// plus kernel-granularity entry points used to generate known answers for the CUDA tests.
#include <atomic>
#include <cmath>
#include <cstring>
#include <string>
#include <thread>

#include "../jbonsai_b200/csrc/host/frontend.h"
#include "oracle.h"

namespace {

struct OrcEngine {
  jb::VoiceSet voices;
  jb::Condition condition;
  explicit OrcEngine(jb::VoiceSet v) : voices(std::move(v)) { condition.load_model(voices); }
};

thread_local std::string g_err;

int fail(const std::exception& e) {
  g_err = e.what();
  if (auto* je = dynamic_cast<const jb::Error*>(&e)) return je->code;
  return jb::kErrArgument;
}

std::vector<std::vector<double>> window_rows(const jb::Windows& w) {
  std::vector<std::vector<double>> r;
  for (const auto& x : w.windows) r.push_back(x.coefficients);
  return r;
}

// MLPG for every stream of one utterance: returns [T][dim] per stream.
void run_mlpg(const OrcEngine& e, const jb::UtterancePack& pack, std::vector<std::vector<double>>* pars,
              size_t* T) {
  pars->clear();
  *T = pack.nframes;
  for (size_t s = 0; s < pack.streams.size(); ++s) {
    const jb::StreamPack& sp = pack.streams[s];
    oracle::MlpgInput in;
    in.nstates = pack.nstates;
    in.durations = pack.durations.data();
    in.dim = sp.dim;
    in.windows = window_rows(e.voices.stream_windows(s));
    in.mean = sp.mean.data();
    in.vari = sp.vari.data();
    in.msd = sp.msd.data();
    in.msd_threshold = e.condition.msd_threshold[s];
    in.has_gv = sp.has_gv;
    in.gv_mean = sp.gv_mean.data();
    in.gv_vari = sp.gv_vari.data();
    in.gv_switch = sp.gv_switch.data();
    in.gv_weight = e.condition.gv_weight[s];
    size_t nf = 0;
    pars->push_back(oracle::mlpg_create(in, &nf));
  }
}

std::vector<double> synth(const OrcEngine& e, const std::vector<std::string>& lines, bool fma,
                          std::vector<double>* exc_tap = nullptr, std::vector<double>* pulse_tap = nullptr,
                          std::vector<double>* gained_tap = nullptr) {
  jb::UtterancePack pack = jb::build_utterance(e.voices, e.condition, lines);
  std::vector<std::vector<double>> pars;
  size_t T = 0;
  run_mlpg(e, pack, &pars, &T);
  size_t nlpf = pars.size() > 2 ? pack.streams[2].dim : 0;
  oracle::Vocoder v(pack.streams[0].dim, nlpf, e.condition.stage, e.condition.use_log_gain,
                    e.condition.sampling_frequency, e.condition.alpha, e.condition.beta,
                    e.condition.volume, e.condition.fperiod, fma);
  v.excitation_tap = exc_tap;
  v.pulse_tap = pulse_tap;
  v.gained_tap = gained_tap;
  std::vector<double> nolpf;
  return oracle::generate_all(&v, T, pars[0].data(), pars[1].data(),
                              nlpf ? pars[2].data() : nolpf.data());
}

double* dup(const std::vector<double>& v) {
  double* p = (double*)malloc(v.size() * sizeof(double) + 8);
  if (!v.empty()) memcpy(p, v.data(), v.size() * sizeof(double));
  return p;
}

}  // namespace

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }
void orc_free_buf(void* p) { free(p); }

int orc_load(const char* const* paths, size_t n, void** out) {
  try {
    std::vector<std::shared_ptr<jb::Voice>> voices;
    for (size_t i = 0; i < n; ++i) {
      std::vector<uint8_t> b = jb::read_file(paths[i]);
      voices.push_back(jb::parse_htsvoice(b.data(), b.size()));
    }
    *out = new OrcEngine(jb::VoiceSet(std::move(voices)));
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

void orc_free(void* h) { delete (OrcEngine*)h; }

int orc_set_condition(void* h, const char* key, size_t idx, double value) {
  try {
    jb::condition_set(((OrcEngine*)h)->condition, key, idx, value);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// kind: 0 duration, 1 parameter, 2 gv
int orc_set_weights(void* h, int kind, size_t stream, const double* w, size_t n) {
  try {
    auto& iw = ((OrcEngine*)h)->condition.interporation_weight;
    std::vector<double> v(w, w + n);
    if (kind == 0) iw.set_duration(v);
    else if (kind == 1) iw.set_parameter(stream, v);
    else iw.set_gv(stream, v);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// Engine::synthesize on the CPU. *out is malloc'ed (orc_free_buf).
int orc_synthesize(void* h, const char* const* lines, size_t n, int fma, double** out, size_t* len) {
  try {
    std::vector<std::string> l(lines, lines + n);
    std::vector<double> w = synth(*(OrcEngine*)h, l, fma != 0);
    *len = w.size();
    *out = dup(w);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// Waveform plus the pre-gain excitation, the pulse train and the gained excitation (MLSA input).
int orc_synthesize_tap(void* h, const char* const* lines, size_t n, int fma, double** out,
                       double** exc, double** pulse, double** gained, size_t* len) {
  try {
    std::vector<std::string> l(lines, lines + n);
    std::vector<double> tap, ptap, gtap;
    std::vector<double> w = synth(*(OrcEngine*)h, l, fma != 0, &tap, &ptap, &gtap);
    *len = w.size();
    *out = dup(w);
    *exc = dup(tap);
    *pulse = dup(ptap);
    *gained = dup(gtap);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// MlpgAdjust::create for streams 0..2: [T][35], [T][1], [T][31] (malloc'ed).
int orc_parameters(void* h, const char* const* lines, size_t n, double** mcp, double** lf0,
                   double** lpf, size_t* nframes) {
  try {
    OrcEngine& e = *(OrcEngine*)h;
    std::vector<std::string> l(lines, lines + n);
    jb::UtterancePack pack = jb::build_utterance(e.voices, e.condition, l);
    std::vector<std::vector<double>> pars;
    size_t T = 0;
    run_mlpg(e, pack, &pars, &T);
    *nframes = T;
    *mcp = dup(pars[0]);
    *lf0 = dup(pars[1]);
    *lpf = dup(pars.size() > 2 ? pars[2] : std::vector<double>());
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// Batch: utterance u owns lines[line_off[u] .. line_off[u+1]); one utterance per thread over
// `nthreads` threads (the reference is single-threaded per utterance; this is how the CPU
// baseline uses all host cores). out_off[u] = first sample of utterance u in *out.
int orc_synthesize_batch(void* h, const char* const* lines, const size_t* line_off, size_t n_utt,
                         int fma, int nthreads, double** out, size_t* out_off) {
  try {
    const OrcEngine& e = *(OrcEngine*)h;
    std::vector<std::vector<double>> waves(n_utt);
    std::atomic<size_t> next{0};
    std::atomic<int> err{0};
    std::string err_msg;
    auto worker = [&]() {
      for (;;) {
        size_t u = next.fetch_add(1);
        if (u >= n_utt) return;
        try {
          std::vector<std::string> l(lines + line_off[u], lines + line_off[u + 1]);
          waves[u] = synth(e, l, fma != 0);
        } catch (const std::exception& ex) {
          if (!err.exchange(1)) err_msg = ex.what();
        }
      }
    };
    if (nthreads < 1) nthreads = 1;
    std::vector<std::thread> th;
    for (int i = 0; i < nthreads; ++i) th.emplace_back(worker);
    for (auto& t : th) t.join();
    if (err.load()) throw jb::Error(jb::kErrArgument, err_msg);
    size_t total = 0;
    for (size_t u = 0; u < n_utt; ++u) {
      out_off[u] = total;
      total += waves[u].size();
    }
    out_off[n_utt] = total;
    double* buf = (double*)malloc(total * sizeof(double) + 8);
    for (size_t u = 0; u < n_utt; ++u)
      if (!waves[u].empty()) memcpy(buf + out_off[u], waves[u].data(), waves[u].size() * sizeof(double));
    *out = buf;
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// ---- kernel-granularity known answers
// First n Gaussian noise samples of the global sequence (excitation.rs:197-231).
void orc_noise(size_t n, double* out) {
  oracle::Random r;
  for (size_t i = 0; i < n; ++i) out[i] = r.nrandom();
}

// Pitch period of a frame, rate / exp(clamp(lf0)) with the host libm (vocoder/mod.rs:73-77): what Vocoder::synthesize
// of this oracle computes per frame.
void orc_pitch_period(const double* lf0, size_t n, double rate, double* out) {
  for (size_t i = 0; i < n; ++i) {
    double p = 0.0;
    if (lf0[i] != -1e10) {
      const double l = lf0[i] < jb::kMinLf0 ? jb::kMinLf0 : (lf0[i] > jb::kMaxLf0 ? jb::kMaxLf0 : lf0[i]);
      p = rate / std::exp(l);
    }
    out[i] = p;
  }
}

// SpeechGenerator over caller-provided [T][dim] parameters.
int orc_vocoder(size_t nframes, size_t nmcp, size_t nlpf, const double* mcp, const double* lf0,
                const double* lpf, size_t rate, size_t fperiod, double alpha, double beta,
                double volume, int fma, double* out, double* exc_or_null) {
  try {
    oracle::Vocoder v(nmcp, nlpf, 0, false, rate, alpha, beta, volume, fperiod, fma != 0);
    std::vector<double> tap;
    if (exc_or_null) v.excitation_tap = &tap;
    std::vector<double> w = oracle::generate_all(&v, nframes, mcp, lf0, lpf);
    memcpy(out, w.data(), w.size() * sizeof(double));
    if (exc_or_null) memcpy(exc_or_null, tap.data(), tap.size() * sizeof(double));
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// The same with Vocoder::new's stage / use_log_gain arguments (stage > 0: `mcp` holds LSP frames).
int orc_vocoder_ex(size_t nframes, size_t nmcp, size_t nlpf, size_t stage, int use_log_gain, const double* mcp,
                   const double* lf0, const double* lpf, size_t rate, size_t fperiod, double alpha, double beta,
                   double volume, int fma, double* out, double* exc_or_null) {
  try {
    oracle::Vocoder v(nmcp, nlpf, stage, use_log_gain != 0, rate, alpha, beta, volume, fperiod, fma != 0);
    std::vector<double> tap;
    if (exc_or_null) v.excitation_tap = &tap;
    std::vector<double> w = oracle::generate_all(&v, nframes, mcp, lf0, lpf);
    memcpy(out, w.data(), w.size() * sizeof(double));
    if (exc_or_null) memcpy(exc_or_null, tap.data(), tap.size() * sizeof(double));
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// Stage::NonZero frame transform: one LSP frame [n] -> coefficients [n] (vocoder/mod.rs:92-107, 148-155).
int orc_lsp_coefficients(const double* spectrum, size_t n, double alpha, int use_log_gain, size_t stage,
                         double beta, int filtered, double* out) {
  try {
    std::vector<double> c = oracle::lsp_frame_coefficients(spectrum, n, alpha, use_log_gain != 0, stage, beta,
                                                           filtered != 0);
    memcpy(out, c.data(), c.size() * sizeof(double));
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// MlpgAdjust::create over caller-provided state-level parameters. windows: nwin rows, row w has
// win_len[w] coefficients, concatenated in win_coef. out: [T][dim], T = sum(durations).
int orc_mlpg(size_t nstates, const int32_t* durations, size_t dim, size_t nwin,
             const size_t* win_len, const double* win_coef, const double* mean, const double* vari,
             const double* msd, double msd_threshold, int has_gv, const double* gv_mean,
             const double* gv_vari, const uint8_t* gv_switch, double gv_weight, double* out) {
  try {
    oracle::MlpgInput in;
    in.nstates = nstates;
    in.durations = durations;
    in.dim = dim;
    const double* c = win_coef;
    for (size_t w = 0; w < nwin; ++w) {
      in.windows.emplace_back(c, c + win_len[w]);
      c += win_len[w];
    }
    in.mean = mean;
    in.vari = vari;
    in.msd = msd;
    in.msd_threshold = msd_threshold;
    in.has_gv = has_gv != 0;
    in.gv_mean = gv_mean;
    in.gv_vari = gv_vari;
    in.gv_switch = gv_switch;
    in.gv_weight = gv_weight;
    size_t T = 0;
    std::vector<double> r = oracle::mlpg_create(in, &T);
    memcpy(out, r.data(), r.size() * sizeof(double));
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// MelCepstrum::postfilter_mcp in place on c[n]; also the impulse-response energies b2en(mc2b(c))
// before and after (the filter's defining property: they are equal).
void orc_postfilter(double* c, size_t n, double alpha, double beta, double* e_before, double* e_after) {
  std::vector<double> v(c, c + n);
  *e_before = oracle::b2en(oracle::mc2b(v.data(), n, alpha), alpha);
  oracle::postfilter_mcp(&v, alpha, beta);
  *e_after = oracle::b2en(oracle::mc2b(v.data(), n, alpha), alpha);
  for (size_t i = 0; i < n; ++i) c[i] = v[i];
}

void orc_boundary_distances(const uint8_t* mask, size_t n, size_t* left, size_t* right) {
  std::vector<uint8_t> m(mask, mask + n);
  std::vector<size_t> l, r;
  oracle::boundary_distances(m, &l, &r);
  for (size_t i = 0; i < n; ++i) {
    left[i] = l[i];
    right[i] = r[i];
  }
}

}  // extern "C"
