// extern "C" layer over jb::Engine (declarations + reference citations: include/jbonsai_b200.h).
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/jbonsai_b200.h"
#include "cuda/kernels.h"
#include "engine.h"

// Streams and resident batches keep their engine alive (handles may be released in any order).
struct jb_engine {
  std::shared_ptr<jb::Engine> e;
};
struct jb_stream {
  std::shared_ptr<jb::Engine> e;
  std::unique_ptr<jb::Stream> s;
};
struct jb_resident {
  std::shared_ptr<jb::Engine> e;
  jb::Engine::Resident* r;
};

namespace {

thread_local std::string g_err;

int fail(const std::exception& ex) {
  g_err = ex.what();
  if (auto* je = dynamic_cast<const jb::Error*>(&ex)) return je->code;
  return JB_ERR_ARGUMENT;
}

#define JB_TRY try {
#define JB_CATCH                     \
  }                                  \
  catch (const std::exception& ex) { \
    return fail(ex);                 \
  }                                  \
  return JB_OK;

std::vector<std::vector<std::string>> to_utts(const char* const* lines, const size_t* line_off,
                                              size_t n_utt) {
  std::vector<std::vector<std::string>> utts(n_utt);
  for (size_t u = 0; u < n_utt; ++u)
    if (line_off[u + 1] < line_off[u]) throw jb::Error(jb::kErrArgument, "line_off must be non-decreasing");
  auto copy_range = [&](size_t a, size_t b) {
    for (size_t u = a; u < b; ++u) utts[u].assign(lines + line_off[u], lines + line_off[u + 1]);
  };
  // ~40 label strings per utterance: for a big batch the copies are worth a few threads
  const size_t nt = n_utt >= 256 ? std::min<size_t>(8, std::max(1u, std::thread::hardware_concurrency())) : 1;  // plain copies
  if (nt <= 1) {
    copy_range(0, n_utt);
  } else {
    std::vector<std::thread> th;
    for (size_t t = 0; t < nt; ++t) th.emplace_back(copy_range, n_utt * t / nt, n_utt * (t + 1) / nt);
    for (auto& t : th) t.join();
  }
  return utts;
}

void fill_timings(const jb::BatchTimings& t, double* out) {
  if (!out) return;
  const double v[20] = {t.host_frontend_ms, t.pack_ms, t.h2d_ms, t.kernels_ms, t.d2h_ms, t.mlpg_ms,
                        t.excitation_ms, t.mlsa_ms, (double)t.h2d_bytes, (double)t.d2h_bytes,
                        (double)t.total_frames, (double)t.total_samples, (double)t.voiced_samples,
                        (double)t.launches, (double)t.mlsa_segments, (double)t.mlsa_seg_frames,
                        t.mlsa_cut_worst, (double)t.mlsa_fallback, (double)t.mlsa_repaired, 0};
  memcpy(out, v, sizeof v);
}

double* dup_vec(const std::vector<double>& v) {
  double* p = (double*)malloc(v.size() * 8 + 8);
  if (!v.empty()) memcpy(p, v.data(), v.size() * 8);
  return p;
}

}  // namespace

extern "C" {

const char* jb_last_error(void) { return g_err.c_str(); }
const char* jb_version(void) { return "jbonsai_b200 0.1.0 (reference jbonsai 0.4.2)"; }

int jb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int jb_device_measure_peaks(int device, double* fp64_fma_tflops, double* hbm_copy_gbs) {
  JB_TRY
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = (cudaError_t)jbk::measure_peaks(fp64_fma_tflops, hbm_copy_gbs);
  if (e != cudaSuccess) throw jb::Error(jb::kErrCuda, std::string("CUDA error: ") + cudaGetErrorString(e));
  JB_CATCH
}

int jb_device_pitch_period(int device, const double* lf0, size_t n, double rate, double* out) {
  JB_TRY
  auto ck = [](cudaError_t e) {
    if (e != cudaSuccess) throw jb::Error(jb::kErrCuda, std::string("CUDA error: ") + cudaGetErrorString(e));
  };
  ck(cudaSetDevice(device));
  double *d_in = nullptr, *d_out = nullptr;
  if (n == 0) return JB_OK;
  ck(cudaMalloc((void**)&d_in, n * 8));
  if (cudaMalloc((void**)&d_out, n * 8) != cudaSuccess) {
    cudaFree(d_in);
    throw jb::Error(jb::kErrCuda, "cudaMalloc failed");
  }
  cudaError_t e = cudaMemcpy(d_in, lf0, n * 8, cudaMemcpyHostToDevice);
  jbk::BatchLayout bl{};
  jbk::VocoderParams vp{};
  vp.rate = rate;
  jbk::VocoderDev vd{};
  vd.lf0 = d_in;
  vd.pitch = d_out;
  if (e == cudaSuccess) e = (cudaError_t)jbk::launch_frame_prep(bl, vp, vd, (int64_t)n, 1, nullptr);  // the kernel the engine runs
  if (e == cudaSuccess) e = cudaMemcpy(out, d_out, n * 8, cudaMemcpyDeviceToHost);
  cudaFree(d_in);
  cudaFree(d_out);
  ck(e);
  JB_CATCH
}

int jb_engine_load(const char* const* paths, size_t n, jb_engine** out) {
  JB_TRY
  std::vector<std::string> p(paths, paths + n);
  *out = new jb_engine{std::shared_ptr<jb::Engine>(jb::Engine::load(p))};
  JB_CATCH
}

int jb_engine_load_from_bytes(const uint8_t* const* bufs, const size_t* lens, size_t n, jb_engine** out) {
  JB_TRY
  std::vector<std::vector<uint8_t>> b(n);
  for (size_t i = 0; i < n; ++i) b[i].assign(bufs[i], bufs[i] + lens[i]);
  *out = new jb_engine{std::shared_ptr<jb::Engine>(jb::Engine::load_from_bytes(b))};
  JB_CATCH
}

void jb_engine_free(jb_engine* e) { delete e; }

int jb_engine_set_condition(jb_engine* e, const char* key, size_t idx, double value) {
  JB_TRY
  jb::condition_set(e->e->condition, key, idx, value);
  JB_CATCH
}

int jb_engine_get_condition(const jb_engine* e, const char* key, size_t idx, double* value) {
  JB_TRY
  *value = jb::condition_get(e->e->condition, key, idx);
  JB_CATCH
}

int jb_engine_set_interpolation_weight(jb_engine* e, int kind, size_t stream, const double* w, size_t n) {
  JB_TRY
  std::vector<double> v(w, w + n);
  auto& iw = e->e->condition.interporation_weight;
  if (kind == 0) iw.set_duration(v);
  else if (kind == 1) iw.set_parameter(stream, v);
  else if (kind == 2) iw.set_gv(stream, v);
  else throw jb::Error(jb::kErrArgument, "kind must be 0, 1 or 2");
  JB_CATCH
}

int jb_engine_set_devices(jb_engine* e, const int* devices, size_t n) {
  JB_TRY
  e->e->set_devices(std::vector<int>(devices, devices + n));
  JB_CATCH
}

int jb_engine_set_mlsa_variant(jb_engine* e, int variant) {
  e->e->mlsa_variant = variant;
  return JB_OK;
}

int jb_engine_set_device_gather(jb_engine* e, int enabled) {
  e->e->device_gather = enabled != 0;
  return JB_OK;
}

int jb_engine_set_output_mode(jb_engine* e, int mode) {
  JB_TRY
  // mode 1 with another number of filter launches: 10 + waves (e.g. 14 = four waves)
  if (mode >= 11 && mode <= 26) {
    e->e->output_mode = 1;
    e->e->output_fractions.assign((size_t)(mode - 10), 1.0);  // equal shares
    return JB_OK;
  }
  if (mode < 0 || mode > 2) throw jb::Error(jb::kErrArgument, "output mode must be 0, 1, 2 or 10 + waves");
  e->e->output_mode = mode;
  if (mode == 1) e->e->output_fractions = jb::Engine::default_output_fractions();
  JB_CATCH
}

int jb_engine_set_output_fractions(jb_engine* e, const double* fractions, size_t n) {
  JB_TRY
  if (n < 1 || n > 16) throw jb::Error(jb::kErrArgument, "1..16 output fractions");
  for (size_t i = 0; i < n; ++i)
    if (!(fractions[i] > 0.0)) throw jb::Error(jb::kErrArgument, "output fractions must be positive");
  e->e->output_mode = 1;
  e->e->output_fractions.assign(fractions, fractions + n);
  JB_CATCH
}

int jb_engine_set_pipeline_chunks(jb_engine* e, int chunks) {
  JB_TRY
  if (chunks < 1 || chunks > 4) throw jb::Error(jb::kErrArgument, "pipeline chunks must be 1..4");
  e->e->pipeline_chunks = chunks;
  JB_CATCH
}

int jb_engine_set_mlsa_segmenting(jb_engine* e, int warm_frames, double cut_tolerance) {
  JB_TRY
  if (warm_frames < 0 || !(cut_tolerance >= 0.0))
    throw jb::Error(jb::kErrArgument, "warm_frames and cut_tolerance must be non-negative");
  e->e->mlsa_warm_frames = warm_frames;
  e->e->mlsa_cut_tolerance = cut_tolerance;
  JB_CATCH
}

int jb_engine_set_mlsa_loud_warm(jb_engine* e, int loud_frames, double threshold) {
  JB_TRY
  if (loud_frames < 0) throw jb::Error(jb::kErrArgument, "loud_frames must be non-negative");
  e->e->mlsa_warm_loud = loud_frames;
  e->e->mlsa_loud_threshold = threshold;
  JB_CATCH
}

int jb_engine_synthesize_batch(const jb_engine* e, const char* const* lines, const size_t* line_off,
                               size_t n_utt, double** out, size_t* out_off, double* timings) {
  JB_TRY
  std::vector<size_t> off;
  jb::BatchTimings tm;
  e->e->synthesize_batch(to_utts(lines, line_off, n_utt), out, &off, nullptr, &tm);
  memcpy(out_off, off.data(), off.size() * sizeof(size_t));
  fill_timings(tm, timings);
  JB_CATCH
}

int jb_engine_synthesize_batch_pcm16(const jb_engine* e, const char* const* lines, const size_t* line_off,
                                     size_t n_utt, int16_t** out, size_t* out_off, double* timings) {
  JB_TRY
  std::vector<size_t> off;
  jb::BatchTimings tm;
  e->e->synthesize_batch_pcm16(to_utts(lines, line_off, n_utt), out, &off, &tm);
  memcpy(out_off, off.data(), off.size() * sizeof(size_t));
  fill_timings(tm, timings);
  JB_CATCH
}

int jb_engine_synthesize(const jb_engine* e, const char* const* lines, size_t n_lines, double** out,
                         size_t* out_len) {
  size_t line_off[2] = {0, n_lines}, out_off[2] = {0, 0};
  int rc = jb_engine_synthesize_batch(e, lines, line_off, 1, out, out_off, nullptr);
  if (rc == JB_OK) *out_len = out_off[1];
  return rc;
}

int jb_alloc_pinned(size_t bytes, void** out) {
  JB_TRY
  if (!out) throw jb::Error(jb::kErrArgument, "alloc_pinned: null argument");
  *out = jb::Engine::alloc_output(bytes);
  JB_CATCH
}

void jb_free(void* p) {
  // pool-owned pinned buffers are recycled; anything else came from malloc
  if (!p) return;
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeHost) {
    jb::Engine::free_output(p);
    return;
  }
  cudaGetLastError();
  free(p);
}

int jb_engine_debug_batch(const jb_engine* e, const char* const* lines, const size_t* line_off,
                          size_t n_utt, double** mcp, double** lf0, double** lpf, double** pulse,
                          double** exc, int64_t* frame_off) {
  JB_TRY
  jb::BatchDebug dbg;
  dbg.want_parameters = true;
  dbg.want_pulse = true;
  dbg.stop_after_excitation = true;
  std::vector<size_t> off;
  double* w = nullptr;
  e->e->synthesize_batch(to_utts(lines, line_off, n_utt), &w, &off, &dbg, nullptr);
  std::vector<double> ex(w, w + off.back());
  jb::Engine::free_output(w);
  *mcp = dup_vec(dbg.mcp);
  *lf0 = dup_vec(dbg.lf0);
  *lpf = dup_vec(dbg.lpf);
  *pulse = dup_vec(dbg.pulse);
  *exc = dup_vec(ex);
  for (size_t u = 0; u <= n_utt; ++u) frame_off[u] = u < dbg.frame_off.size() ? dbg.frame_off[u] : 0;
  JB_CATCH
}

int jb_engine_noise(const jb_engine* e, size_t n, double* out) {
  JB_TRY
  std::vector<double> v = e->e->noise_prefix(n);
  memcpy(out, v.data(), n * 8);
  JB_CATCH
}

int jb_engine_durations(const jb_engine* e, const char* const* lines, size_t n_lines,
                        int32_t* durations, size_t cap, size_t* n_states) {
  JB_TRY
  std::vector<std::string> l(lines, lines + n_lines);
  jb::UtterancePack p = jb::build_utterance(e->e->voices(), e->e->condition, l);
  *n_states = p.nstates;
  if (p.nstates > cap) throw jb::Error(jb::kErrArgument, "durations buffer too small");
  if (p.nstates) memcpy(durations, p.durations.data(), p.nstates * 4);
  JB_CATCH
}

int jb_engine_tree_index(const jb_engine* e, size_t voice, int model, size_t stream,
                         size_t state_index, const char* label, long* tree_index, long* pdf_index) {
  JB_TRY
  const jb::VoiceSet& vs = e->e->voices();
  if (voice >= vs.size()) throw jb::Error(jb::kErrArgument, "voice index out of range");
  const jb::Voice& v = *vs.voices[voice];
  const jb::Model* m = nullptr;
  if (model == 0) m = &v.duration_model;
  else {
    if (stream >= v.stream_models.size()) throw jb::Error(jb::kErrArgument, "stream out of range");
    if (model == 1) m = &v.stream_models[stream].stream_model;
    else if (v.stream_models[stream].has_gv) m = &v.stream_models[stream].gv_model;
    else throw jb::Error(jb::kErrArgument, "stream has no GV model");
  }
  m->get_index(state_index, label, tree_index, pdf_index);
  JB_CATCH
}

int jb_engine_stream_parameters(const jb_engine* e, const char* const* lines, size_t n_lines,
                                size_t stream, double* mean, double* vari, double* msd, size_t cap,
                                size_t* n_states, size_t* width) {
  JB_TRY
  std::vector<std::string> l(lines, lines + n_lines);
  jb::UtterancePack p = jb::build_utterance(e->e->voices(), e->e->condition, l);
  if (stream >= p.streams.size()) throw jb::Error(jb::kErrArgument, "stream out of range");
  const jb::StreamPack& sp = p.streams[stream];
  *n_states = p.nstates;
  *width = sp.dim * sp.nwin;
  if (sp.mean.size() > cap) throw jb::Error(jb::kErrArgument, "parameter buffer too small");
  if (!sp.mean.empty()) {
    memcpy(mean, sp.mean.data(), sp.mean.size() * 8);
    memcpy(vari, sp.vari.data(), sp.vari.size() * 8);
    memcpy(msd, sp.msd.data(), sp.msd.size() * 8);
  }
  JB_CATCH
}

int jb_engine_duration_parameters(const jb_engine* e, const char* const* lines, size_t n_lines,
                                  double* mean, double* vari, size_t cap, size_t* n_states) {
  JB_TRY
  std::vector<std::string> l(lines, lines + n_lines);
  const jb::Condition& c = e->e->condition;
  jb::Labels labels = jb::Labels::load_from_strings(c.sampling_frequency, c.fperiod, l);
  jb::Models models{labels.labels, e->e->voices(), c.interporation_weight};
  std::vector<jb::MeanVari> d = models.duration();
  *n_states = d.size();
  if (d.size() > cap) throw jb::Error(jb::kErrArgument, "duration buffer too small");
  for (size_t i = 0; i < d.size(); ++i) {
    mean[i] = d[i].mean;
    vari[i] = d[i].vari;
  }
  JB_CATCH
}

int jb_engine_gv_parameters(const jb_engine* e, const char* const* lines, size_t n_lines,
                            size_t stream, double* mean, double* vari, uint8_t* gv_switch,
                            size_t cap, size_t* dim, size_t* n_states, int* has_gv) {
  JB_TRY
  std::vector<std::string> l(lines, lines + n_lines);
  jb::UtterancePack p = jb::build_utterance(e->e->voices(), e->e->condition, l);
  if (stream >= p.streams.size()) throw jb::Error(jb::kErrArgument, "stream out of range");
  const jb::StreamPack& sp = p.streams[stream];
  *has_gv = sp.has_gv;
  *dim = sp.dim;
  *n_states = p.nstates;
  if (sp.has_gv) {
    if (sp.dim > cap || p.nstates > cap) throw jb::Error(jb::kErrArgument, "gv buffer too small");
    memcpy(mean, sp.gv_mean.data(), sp.dim * 8);
    memcpy(vari, sp.gv_vari.data(), sp.dim * 8);
    memcpy(gv_switch, sp.gv_switch.data(), p.nstates);
  }
  JB_CATCH
}

int jb_resident_prepare(const jb_engine* e, const char* const* lines, const size_t* line_off,
                        size_t n_utt, int device_index, jb_resident** out) {
  JB_TRY
  *out = new jb_resident{e->e, e->e->prepare(to_utts(lines, line_off, n_utt), device_index)};
  JB_CATCH
}

static jb::Engine::VocoderConfig to_cfg(const jb_vocoder_config* c) {
  if (!c) throw jb::Error(jb::kErrArgument, "null vocoder config");
  jb::Engine::VocoderConfig v;
  v.nmcp = c->nmcp;
  v.nlpf = c->nlpf;
  v.stage = c->stage;
  v.use_log_gain = c->use_log_gain != 0;
  v.rate = c->rate;
  v.alpha = c->alpha;
  v.beta = c->beta;
  v.volume = c->volume;
  v.fperiod = c->fperiod;
  v.coefficient_rows = c->coefficient_rows != 0;
  return v;
}

int jb_resident_prepare_tracks(const jb_engine* e, const jb_vocoder_config* cfg, size_t n_utt, const size_t* frame_off,
                               const double* spectrum, const double* lf0, const double* lpf, int device_index,
                               jb_resident** out) {
  JB_TRY
  *out = new jb_resident{e->e, e->e->prepare_tracks(to_cfg(cfg), n_utt, frame_off, spectrum, lf0, lpf, device_index)};
  JB_CATCH
}

int jb_engine_vocoder_synthesize(const jb_engine* e, const jb_vocoder_config* cfg, size_t n_utt,
                                 const size_t* frame_off, const double* spectrum, const double* lf0,
                                 const double* lpf, double** out, size_t* out_off) {
  JB_TRY
  jb::Engine::Resident* r = e->e->prepare_tracks(to_cfg(cfg), n_utt, frame_off, spectrum, lf0, lpf, 0);
  try {
    e->e->run_resident(r, nullptr);
    std::vector<size_t> off;
    e->e->fetch_resident(r, out, &off);
    if (out_off) memcpy(out_off, off.data(), off.size() * sizeof(size_t));
  } catch (...) {
    e->e->release(r);
    throw;
  }
  e->e->release(r);
  JB_CATCH
}

int jb_resident_rows(const jb_engine* e, jb_resident* r, double* out, size_t cap) {
  JB_TRY
  std::vector<double> v = r->e->resident_rows(r->r);
  if (v.size() > cap) throw jb::Error(jb::kErrArgument, "jb_resident_rows: buffer too small");
  if (!v.empty()) memcpy(out, v.data(), v.size() * sizeof(double));
  JB_CATCH
}

int jb_resident_run(const jb_engine* e, jb_resident* r, double* device_ms, double* timings) {
  JB_TRY
  jb::BatchTimings tm;
  double ms = r->e->run_resident(r->r, &tm);
  if (device_ms) *device_ms = ms;
  fill_timings(tm, timings);
  JB_CATCH
}

int jb_resident_fetch(const jb_engine* e, jb_resident* r, double** out, size_t* out_off) {
  JB_TRY
  std::vector<size_t> off;
  r->e->fetch_resident(r->r, out, &off);
  memcpy(out_off, off.data(), off.size() * sizeof(size_t));
  JB_CATCH
}

int jb_resident_device_output(const jb_engine* e, jb_resident* r, const double** d_out, size_t* out_off, int* device) {
  JB_TRY
  std::vector<size_t> off;
  r->e->resident_device_output(r->r, d_out, &off, device);
  if (out_off) memcpy(out_off, off.data(), off.size() * sizeof(size_t));
  JB_CATCH
}

int jb_engine_synthesize_batch_device(const jb_engine* e, const char* const* lines, const size_t* line_off, size_t n_utt,
                                      int device_index, jb_resident** keep, const double** d_out, size_t* out_off,
                                      double* timings) {
  JB_TRY
  jb::Engine::Resident* r = e->e->prepare(to_utts(lines, line_off, n_utt), device_index);
  try {
    jb::BatchTimings tm;
    e->e->run_resident(r, &tm);
    std::vector<size_t> off;
    e->e->resident_device_output(r, d_out, &off, nullptr);
    if (out_off) memcpy(out_off, off.data(), off.size() * sizeof(size_t));
    fill_timings(tm, timings);
  } catch (...) {
    e->e->release(r);
    throw;
  }
  *keep = new jb_resident{e->e, r};
  JB_CATCH
}

int jb_engine_mlpg_create(const jb_engine* e, const jb_mlpg_input* in, int device_index, double** out, size_t* frame_off) {
  JB_TRY
  if (!in || !out) throw jb::Error(jb::kErrArgument, "mlpg: null argument");
  jb::Engine::MlpgInput m;
  m.n_utt = in->n_utt;
  m.dim = in->dim;
  m.state_off = in->state_off;
  m.durations = in->durations;
  if (in->n_windows > 8) throw jb::Error(jb::kErrUnsupported, "more than 4 windows per stream");
  size_t o = 0;
  for (size_t w = 0; w < in->n_windows; ++w) {
    m.windows.emplace_back(in->window_coefficients + o, in->window_coefficients + o + in->window_widths[w]);
    o += in->window_widths[w];
  }
  m.mean = in->mean;
  m.vari = in->vari;
  m.msd = in->msd;
  m.msd_threshold = in->msd_threshold;
  m.gv_mean = in->gv_mean;
  m.gv_vari = in->gv_vari;
  m.gv_switch = in->gv_switch;
  m.gv_weight = in->gv_weight;
  std::vector<double> par;
  std::vector<size_t> foff;
  e->e->mlpg_create(m, device_index, &par, &foff);
  *out = dup_vec(par);
  if (frame_off) memcpy(frame_off, foff.data(), foff.size() * sizeof(size_t));
  JB_CATCH
}

// No exceptions and no containers on this path: nothing here may allocate on the device, and the host side keeps
// to static strings as well.
int jb_mlpg_solve_noalloc(int device, const double* d_mean, const double* d_vari, size_t n_frames, size_t dim,
                          size_t n_windows, const size_t* window_widths, const double* window_coefficients,
                          double* d_out, void* stream) {
  static_assert(JB_MLPG_NOALLOC_MAX_CELLS == jbk::kMlpgNoAllocMaxCells, "header and kernel limit differ");
  auto bad = [](const char* msg) {
    g_err = msg;
    return (int)JB_ERR_ARGUMENT;
  };
  if (dim == 0 || n_windows == 0 || !window_widths || !window_coefficients) return bad("mlpg_solve_noalloc: dim, n_windows and the windows must be given");
  if (n_windows > (size_t)jbk::kMaxWin) return bad("mlpg_solve_noalloc: more than 4 windows");
  jbk::WindowSet ws{};
  ws.nwin = (int)n_windows;
  size_t o = 0;
  for (size_t w = 0; w < n_windows; ++w) {
    const size_t width = window_widths[w];
    if (width == 0 || width % 2 == 0 || width > (size_t)jbk::kMaxWinWidth) return bad("mlpg_solve_noalloc: window widths must be odd and <= 5");
    ws.width[w] = (int)width;
    for (size_t k = 0; k < width; ++k) ws.coef[w][k] = window_coefficients[o + k];
    o += width;
  }
  if (n_frames == 0) return JB_OK;
  if (dim > (size_t)JB_MLPG_NOALLOC_MAX_CELLS || n_frames > (size_t)JB_MLPG_NOALLOC_MAX_CELLS / dim)
    return bad("mlpg_solve_noalloc: n_frames * dim exceeds JB_MLPG_NOALLOC_MAX_CELLS (131072), the size of the static device scratch");
  if (!d_mean || !d_vari || !d_out) return bad("mlpg_solve_noalloc: null device pointer");
  const char* what = "";
  const int ce = jbk::mlpg_solve_noalloc(device, d_mean, d_vari, (long long)n_frames, (int)dim, ws, d_out,
                                         (cudaStream_t)stream, &what);
  if (ce != 0) {
    g_err = "mlpg_solve_noalloc: CUDA error in ";
    g_err += what;
    g_err += ": ";
    g_err += cudaGetErrorString((cudaError_t)ce);
    return (int)JB_ERR_CUDA;
  }
  return JB_OK;
}

int jb_debug_plan_segments(const int64_t* frames, size_t n_utt, int warm, int warm_loud, double threshold,
                           const double* gain, int64_t target, int32_t* out, size_t cap, size_t* n_segments) {
  JB_TRY
  if (!frames || !n_segments || target <= 0 || warm < 0) throw jb::Error(jb::kErrArgument, "plan_segments: bad argument");
  const std::vector<int32_t> rows =
      jb::Engine::plan_segments_debug(std::vector<int64_t>(frames, frames + n_utt), warm, warm_loud, threshold, gain, target);
  *n_segments = rows.size() / 5;
  if (out && cap >= rows.size()) memcpy(out, rows.data(), rows.size() * sizeof(int32_t));
  JB_CATCH
}

int jb_set_host_threads(int n) {
  jb::Engine::set_host_threads(n);
  return JB_OK;
}

size_t jb_trim_pools(void) { return jb::Engine::trim_pools(); }

void jb_resident_free(const jb_engine* e, jb_resident* r) {
  if (!r) return;
  try {
    r->e->release(r->r);
  } catch (...) {
  }
  delete r;
}

int jb_stream_open(const jb_engine* e, const char* const* lines, const size_t* line_off,
                   size_t n_utt, int device, jb_stream** out) {
  JB_TRY
  *out = new jb_stream{e->e, std::unique_ptr<jb::Stream>(
                                new jb::Stream(e->e.get(), to_utts(lines, line_off, n_utt), device))};
  JB_CATCH
}

size_t jb_stream_fperiod(const jb_stream* s) { return s->s->fperiod(); }
size_t jb_stream_synthesized_frames(const jb_stream* s) { return s->s->synthesized_frames(); }

int jb_stream_step(jb_stream* s, size_t n_frames, double* buf, size_t cap, size_t* written,
                   double* device_ms) {
  JB_TRY
  if (cap < s->s->n_utt() * n_frames * s->s->fperiod())
    throw jb::Error(jb::kErrArgument, "The length of speech buffer must be larger than fperiod.");
  double ms = s->s->step(n_frames, buf, written);
  if (device_ms) *device_ms = ms;
  JB_CATCH
}

void jb_stream_close(jb_stream* s) { delete s; }

}  // extern "C"
