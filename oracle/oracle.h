// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement of jbonsai's hot path (MLPG solve + mel-cepstral vocoder), written from the
// reference's Rust sources, one function per reference function with the file:line it follows.
// Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
// legs may build, link, import or execute anything under `oracle/`. The product
// (`jbonsai_b200/`) never does: it fails loudly when its CUDA library is missing.
//
// Parity status: PINNED. `tests/test_oracle_goldens.py` checks this code end to end (through
// the shared host front-end in jbonsai_b200/csrc/host) against every golden the reference's
// own tests hold for the path: the seven `src/lib.rs:38-160` waveform tests (lengths exact,
// 14 pinned samples at the reference's own eps 1e-10), `src/duration.rs:141-180`,
// `src/model/mod.rs:184-431`, `src/label.rs:166-197`, `src/mlpg_adjust/mask.rs:90-159`.
// The reference itself (Rust) cannot be compiled here: no cargo/rustc in the image.
#pragma once

#include <cstddef>
#include <cstdint>
#include <vector>

namespace oracle {

constexpr double kMaxLf0 = 9.903487552536127;  // src/constants.rs:4-13
constexpr double kMinLf0 = 2.995732273553991;
constexpr double kNoData = -1e10;

// ---- MLPG (src/mlpg_adjust/{mod,mask,mlpg}.rs)
struct MlpgInput {
  size_t nstates = 0;
  const int32_t* durations = nullptr;  // [S]
  size_t dim = 0;                      // vector_length
  std::vector<std::vector<double>> windows;  // window coefficient rows
  const double* mean = nullptr;        // [S][nwin*dim], index dim*w + d   (mod.rs:62)
  const double* vari = nullptr;        // [S][nwin*dim]
  const double* msd = nullptr;         // [S]; DBL_MAX for non-MSD streams (model/mod.rs:113)
  double msd_threshold = 0.5;
  bool has_gv = false;
  const double* gv_mean = nullptr;     // [dim]
  const double* gv_vari = nullptr;     // [dim]
  const uint8_t* gv_switch = nullptr;  // [S]
  double gv_weight = 1.0;
};

// MlpgAdjust::create (mod.rs:51-95): returns [T][dim] row-major, NODATA in masked frames.
std::vector<double> mlpg_create(const MlpgInput& in, size_t* nframes);

// Mask helpers, exposed for the reference's unit tests (mask.rs:51-82).
void boundary_distances(const std::vector<uint8_t>& mask, std::vector<size_t>* left,
                        std::vector<size_t>* right);

// ---- excitation RNG (src/vocoder/excitation.rs:178-237)
struct Random {
  double queue[64];
  double s[32];
  size_t used = 64;
  uint64_t next = 1;
  double nrandom();
  void fill_queue();
};

// ---- vocoder (src/vocoder/{mod,excitation,mlsa,mglsa,lsp,generalized,cepstrum,coefficients}.rs)
// Parity status of the stage != 0 arm (LSP -> MGLSA): UNPINNED. The reference holds no test, golden or
// shipped voice for it (no .htsvoice under models/ sets GAMMA); it is a literal restatement only.
struct Vocoder {
  Vocoder(size_t nmcp, size_t nlpf, size_t stage, bool use_log_gain, size_t rate, double alpha,
          double beta, double volume, size_t fperiod, bool fma_fir);
  // Vocoder::synthesize (mod.rs:72-178): one frame, writes fperiod samples.
  void synthesize(double lf0, const double* spectrum, const double* lpf, double* rawdata);

  // optional taps for tests: per-sample excitation before gain (ring-buffer output)
  std::vector<double>* excitation_tap = nullptr;
  std::vector<double>* pulse_tap = nullptr;   // pulse amplitude (0 when none) per sample
  std::vector<double>* gained_tap = nullptr;  // excitation * exp(c[0]): the MLSA filter input

  size_t nmcp, nlpf, rate, fperiod;
  double alpha, beta, volume;
  bool fma_fir;
  size_t stage = 0;           // 0: MLSA; n > 0: LSP input, MGLSA filter of n sections with gamma = -1/n (stage.rs:28-38)
  bool use_log_gain = false;
  std::vector<std::vector<double>> mglsa_d;  // MelGeneralizedLogSpectrumApproximation::d (mglsa.rs:5-13)
  bool is_first = true;
  std::vector<double> coefficients;  // Stage::Zero coefficients
  // MelLogSpectrumApproximation<6> state (mlsa.rs:5-10)
  double d11[6] = {0}, d12[6] = {0}, d22[6] = {0};
  std::vector<double> d21[6];
  // Excitation state (excitation.rs:2-10)
  double pitch_of_curr_point = 0.0, pitch_counter = 0.0, pitch_inc_per_point = 0.0;
  std::vector<double> ring;
  size_t ring_index = 0;
  Random random;
};

// helpers exposed for kernel-granularity known-answer tests
std::vector<double> mc2b(const double* c, size_t n, double alpha);       // cepstrum.rs:139-149
void postfilter_mcp(std::vector<double>* c, double alpha, double beta);  // cepstrum.rs:23-37
double b2en(const std::vector<double>& b, double alpha);                // coefficients.rs:75-78
double fir(double* d, size_t dlen, double x, double alpha, const double* coefficients,
           bool fma);                                                    // mlsa.rs:128-163
// Stage::NonZero: one LSP frame -> MGLSA filter coefficients (vocoder/mod.rs:92-107 when !filtered,
// :148-155 when filtered: postfilter_lsp + check_lsp_stability first).
std::vector<double> lsp_frame_coefficients(const double* spectrum, size_t n, double alpha, bool use_log_gain,
                                           size_t stage, double beta, bool filtered);

// SpeechGenerator::generate_all (speech.rs:87-96) over [T][dim] parameter arrays.
std::vector<double> generate_all(Vocoder* v, size_t nframes, const double* mcp, const double* lf0,
                                 const double* lpf);

}  // namespace oracle
