// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h). Compile with -ffp-contract=off; the only
// fused multiply-adds are the explicit std::fma calls of the FMA flavour of `fir`.
// Literal restatement of src/vocoder/{mod,excitation,mlsa,cepstrum,coefficients}.rs, speech.rs.
#include <cmath>
#include <stdexcept>

#include "oracle.h"

namespace oracle {

// ---------------------------------------------------------------- RNG (excitation.rs:178-237)
static inline double rnd(uint64_t* next) {
  *next = *next * 1103515245ull + 12345ull;  // usize wrapping arithmetic (64-bit)
  uint64_t r = (*next / 65536ull) % 32768ull;
  return (double)r / 32767.0;
}

void Random::fill_queue() {
  size_t i = 0;
  while (i < 32) {
    double r1 = 2.0 * rnd(&next) - 1.0;
    double r2 = 2.0 * rnd(&next) - 1.0;
    double ss = r1 * r1 + r2 * r2;
    if (0.0 < ss && ss < 1.0) {
      queue[2 * i] = r1;
      queue[2 * i + 1] = r2;
      s[i] = ss;
      i += 1;
    }
  }
  for (size_t k = 0; k < 32; ++k) {
    double m = std::sqrt(-2.0 * std::log(s[k]) / s[k]);
    queue[2 * k] *= m;
    queue[2 * k + 1] *= m;
  }
}

double Random::nrandom() {
  if (used >= 64) {
    fill_queue();
    used = 0;
  }
  return queue[used++];
}

// ---------------------------------------------------------------- cepstrum helpers
// CepstrumT::mc2b (cepstrum.rs:139-149)
std::vector<double> mc2b(const double* c, size_t n, double alpha) {
  std::vector<double> b(c, c + n);
  if (alpha != 0.0 && n > 0) {
    size_t last = n - 1;
    b[last] = c[last];
    for (size_t i = last; i-- > 0;) b[i] = c[i] - alpha * b[i + 1];
  }
  return b;
}

// CoefficientsT::b2mc (coefficients.rs:65-73)
static std::vector<double> b2mc(const std::vector<double>& b, double alpha) {
  std::vector<double> c(b.size(), 0.0);
  size_t last = b.size() - 1;
  c[last] = b[last];
  for (size_t i = last; i-- > 0;) c[i] = b[i] + alpha * b[i + 1];
  return c;
}

// CepstrumT::freqt (cepstrum.rs:153-173)
static std::vector<double> freqt(const std::vector<double>& c1, size_t m2, double alpha) {
  double aa = 1.0 - alpha * alpha;
  std::vector<double> c2(m2 + 1, 0.0), f(m2 + 1, 0.0);
  for (size_t i = 0; i < c1.size(); ++i) {
    f[0] = c2[0];
    c2[0] = c1[i] + alpha * c2[0];
    if (1 <= m2) {
      f[1] = c2[1];
      c2[1] = aa * f[0] + alpha * c2[1];
    }
    for (size_t j = 2; j < c2.size(); ++j) {
      f[j] = c2[j];
      c2[j] = f[j - 1] + alpha * (c2[j] - c2[j - 1]);
    }
  }
  return c2;
}

// CepstrumT::c2ir (cepstrum.rs:175-186)
static std::vector<double> c2ir(const std::vector<double>& c, size_t len) {
  std::vector<double> ir(len, 0.0);
  ir[0] = std::exp(c[0]);
  for (size_t n = 1; n < len; ++n) {
    double d = 0.0;
    size_t kmax = c.size() < n + 1 ? c.size() : n + 1;
    for (size_t k = 1; k < kmax; ++k) d += (double)k * c[k] * ir[n - k];
    ir[n] = d / (double)n;
  }
  return ir;
}

// CoefficientsT::b2en (coefficients.rs:75-78)
double b2en(const std::vector<double>& b, double alpha) {
  std::vector<double> ir = c2ir(freqt(b2mc(b, alpha), 576 - 1, -alpha), 576);
  double s = 0.0;
  for (double x : ir) s += x * x;
  return s;
}

// MelCepstrum::postfilter_mcp (cepstrum.rs:23-37)
void postfilter_mcp(std::vector<double>* c, double alpha, double beta) {
  if (beta > 0.0 && c->size() > 2) {
    std::vector<double> b = mc2b(c->data(), c->size(), alpha);
    double e1 = b2en(b, alpha);
    b[1] -= beta * alpha * b[2];
    for (size_t k = 2; k < c->size(); ++k) b[k] *= 1.0 + beta;
    double e2 = b2en(b, alpha);
    b[0] += std::log(e1 / e2) / 2.0;
    *c = b2mc(b, alpha);
  }
}

// ---------------------------------------------------------------- MLSA (mlsa.rs)
static const double PPADE[6] = {1.00000000000, 0.49993910000, 0.11070980000,
                                0.01369984000, 0.00095648530, 0.00003041721};  // mlsa.rs:31

// fir (mlsa.rs:128-163); `fma` selects the mul_add! expansion of mlsa.rs:82-115.
double fir(double* d, size_t dlen, double x, double alpha, const double* coefficients, bool fma) {
  const double a = alpha, aa = a * a, aaaa = aa * aa, iaa = 1.0 - aa;
  double rem = -a * x + d[1];
  d[0] = a * x;
  d[1] = iaa * x + a * d[1];
  double y0 = 0.0, y1 = 0.0;
  const double* c = coefficients + 2;
  double* dd = d + 2;
  size_t n = dlen - 2, nchunk = n / 4;
  for (size_t q = 0; q < nchunk; ++q, c += 4, dd += 4) {
    const double d0 = dd[0], d1 = dd[1], d2 = dd[2], d3 = dd[3];
    double n0, n1, n2, n3, nrem;
    if (fma) {
      n0 = std::fma(iaa, rem, a * d0);
      n1 = std::fma(iaa, std::fma(-a, rem, d0), a * d1);
      n2 = std::fma(iaa, std::fma(aa, rem, std::fma(-a, d0, d1)), a * d2);
      n3 = std::fma(iaa, std::fma(aa, std::fma(-a, rem, d0), std::fma(-a, d1, d2)), a * d3);
      nrem = std::fma(aaaa, rem, std::fma(aa, std::fma(-a, d0, d1), std::fma(-a, d2, d3)));
    } else {
      n0 = iaa * rem + a * d0;
      n1 = iaa * (-a * rem + d0) + a * d1;
      n2 = iaa * (aa * rem + (-a * d0 + d1)) + a * d2;
      n3 = iaa * (aa * (-a * rem + d0) + (-a * d1 + d2)) + a * d3;
      nrem = aaaa * rem + (aa * (-a * d0 + d1) + (-a * d2 + d3));
    }
    dd[0] = n0;
    dd[1] = n1;
    dd[2] = n2;
    dd[3] = n3;
    rem = nrem;
    y0 += c[0] * n0 + c[2] * n2;
    y1 += c[1] * n1 + c[3] * n3;
  }
  for (size_t k = nchunk * 4; k < n; ++k, ++c, ++dd) {
    const double dv = *dd;
    double nd = iaa * rem + a * dv;
    rem = -a * rem + dv;
    *dd = nd;
    y0 += *c * nd;
  }
  return y0 + y1;
}

// ---------------------------------------------------------------- stage != 0: LSP -> MGLSA coefficients
// Generalized::gnorm / ignorm (generalized.rs:6-38); gamma != 0 on this path (stage.rs:30)
static std::vector<double> gnorm(const std::vector<double>& c, double gamma) {
  std::vector<double> t(c);
  if (gamma != 0.0) {
    double k = 1.0 + gamma * c[0];
    t[0] = std::pow(k, 1.0 / gamma);
    for (size_t i = 1; i < c.size(); ++i) t[i] = c[i] / k;
  } else {
    t[0] = std::exp(c[0]);
  }
  return t;
}
static std::vector<double> ignorm(const std::vector<double>& c, double gamma) {
  std::vector<double> t(c);
  if (gamma != 0.0) {
    double k = std::pow(c[0], gamma);
    t[0] = (k - 1.0) / gamma;
    for (size_t i = 1; i < c.size(); ++i) t[i] = c[i] * k;
  } else {
    t[0] = std::log(c[0]);
  }
  return t;
}

// MelGeneralizedCepstrum::gc2gc (cepstrum.rs:68-92): g1 = gamma of the input, g2 = of the output
static std::vector<double> gc2gc(const std::vector<double>& c1, double g1, size_t m2, double g2) {
  std::vector<double> c2(m2 + 1, 0.0);
  c2[0] = c1[0];
  for (size_t i = 1; i <= m2; ++i) {
    double ss1 = 0.0, ss2 = 0.0;
    size_t kmax = c1.size() < i ? c1.size() : i;
    for (size_t k = 1; k < kmax; ++k) {
      size_t mk = i - k;
      double cc = c1[k] * c2[mk];
      ss1 += (double)mk * cc;
      ss2 += (double)k * cc;
    }
    if (i < c1.size()) c2[i] = c1[i] + (g2 * ss2 - g1 * ss1) / (double)i;
    else c2[i] = (g2 * ss2 - g1 * ss1) / (double)i;
  }
  return c2;
}

// MelGeneralizedCepstrum::mgc2mgc (cepstrum.rs:94-101)
static std::vector<double> mgc2mgc(const std::vector<double>& c, double a1, double g1, size_t m2, double a2,
                                   double g2) {
  if (a1 == a2) return ignorm(gc2gc(gnorm(c, g1), g1, m2, g2), g2);
  double a = (a2 - a1) / (1.0 - a1 * a2);
  return ignorm(gc2gc(gnorm(freqt(c, m2, a), g1), g1, m2, g2), g2);
}

// LineSpectralPairs::lsp2lpc (lsp.rs:27-88). As in the reference, the whole vector (gain slot
// included) is read as m = len line spectral frequencies; the result has m + 1 entries.
static std::vector<double> lsp2lpc(const std::vector<double>& lsp) {
  const size_t m = lsp.size();
  const size_t mh1 = m % 2 == 0 ? m / 2 : (m + 1) / 2, mh2 = m % 2 == 0 ? m / 2 : (m - 1) / 2;
  std::vector<double> p, q;
  for (size_t i = 0; i < m; i += 2) p.push_back(-2.0 * std::cos(lsp[i]));
  for (size_t i = 1; i < m; i += 2) q.push_back(-2.0 * std::cos(lsp[i]));
  std::vector<double> a0(mh1 + 1, 0.0), a1(mh1 + 1, 0.0), a2(mh1 + 1, 0.0);
  std::vector<double> b0(mh2 + 1, 0.0), b1(mh2 + 1, 0.0), b2(mh2 + 1, 0.0);
  double xff = 0.0, xf = 0.0;
  std::vector<double> c(m + 1, 0.0);
  for (size_t k = 0; k <= m; ++k) {
    double xx = k == 0 ? 1.0 : 0.0;
    if (m % 2 == 1) {
      a0[0] = xx;
      b0[0] = xx - xff;
      xff = xf;
      xf = xx;
    } else {
      a0[0] = xx + xf;
      b0[0] = xx - xf;
      xf = xx;
    }
    for (size_t i = 0; i < mh1; ++i) {
      a0[i + 1] = a0[i] + p[i] * a1[i] + a2[i];
      a2[i] = a1[i];
      a1[i] = a0[i];
    }
    for (size_t i = 0; i < mh2; ++i) {
      b0[i + 1] = b0[i] + q[i] * b1[i] + b2[i];
      b2[i] = b1[i];
      b1[i] = b0[i];
    }
    if (k > 0) c[k - 1] = -0.5 * (a0[mh1] + b0[mh2]);
  }
  for (size_t i = m; i-- > 0;) c[i + 1] = -c[i];
  c[0] = 1.0;
  return c;
}

// LineSpectralPairs::lsp2mgc (lsp.rs:90-102)
static std::vector<double> lsp2mgc(const std::vector<double>& lsp, double alpha, bool use_log_gain, size_t stage,
                                   double gamma) {
  std::vector<double> lpc = lsp2lpc(lsp);
  lpc[0] = use_log_gain ? std::exp(lsp[0]) : lsp[0];
  lpc = ignorm(lpc, gamma);
  for (size_t i = 1; i < lpc.size(); ++i) lpc[i] *= -(double)stage;
  return mgc2mgc(lpc, alpha, gamma, lsp.size() - 1, alpha, gamma);
}

// LineSpectralPairs::postfilter_lsp (lsp.rs:108-137)
static void postfilter_lsp(std::vector<double>* l, double alpha, bool use_log_gain, size_t stage, double gamma,
                           double beta) {
  std::vector<double>& s = *l;
  const size_t n = s.size();
  if (!(beta > 0.0) || n <= 2) return;
  auto en = [&]() {
    double e = 0.0;
    for (double x : lsp2mgc(s, alpha, use_log_gain, stage, gamma)) e += x * x;
    return e;
  };
  std::vector<double> buf(n, 0.0);
  const double en1 = en();
  for (size_t i = 0; i < n; ++i) {
    if (i > 1 && i < n - 1) {
      double d1 = beta * (s[i + 1] - s[i]);
      double d2 = beta * (s[i] - s[i - 1]);
      buf[i] = s[i - 1] + d2 + (d2 * d2 * ((s[i + 1] - s[i - 1]) - (d1 + d2))) / ((d2 * d2) + (d1 * d1));
    } else {
      buf[i] = s[i];
    }
  }
  s = buf;
  const double en2 = en();
  if (en1 != en2) {
    if (use_log_gain) s[0] += 0.5 * std::log(en1 / en2);
    else s[0] *= std::sqrt(en1 / en2);
  }
}

// LineSpectralPairs::check_lsp_stability (lsp.rs:139-165)
static void check_lsp_stability(std::vector<double>* l) {
  std::vector<double>& s = *l;
  const double pi = 3.14159265358979323846;
  const double min = 0.25 * pi / (double)s.size();
  const size_t last = s.size() - 1;
  for (int it = 0; it < 4; ++it) {
    bool find = false;
    for (size_t j = 1; j < last; ++j) {
      double tmp = s[j + 1] - s[j];
      if (tmp < min) {
        s[j] -= 0.5 * (min - tmp);
        s[j + 1] += 0.5 * (min - tmp);
        find = true;
      }
    }
    if (s[1] < min) {
      s[1] = min;
      find = true;
    }
    if (s[last] > pi - min) {
      s[last] = pi - min;
      find = true;
    }
    if (!find) break;
  }
}

// The Stage::NonZero frame transform of Vocoder::synthesize (mod.rs:92-107 for the first frame,
// :148-155 for every frame): LSP frame -> filter coefficients.
std::vector<double> lsp_frame_coefficients(const double* spectrum, size_t n, double alpha, bool use_log_gain,
                                           size_t stage, double beta, bool filtered) {
  const double gamma = -1.0 / (double)stage;
  std::vector<double> lsp(spectrum, spectrum + n);
  if (filtered) {
    postfilter_lsp(&lsp, alpha, use_log_gain, stage, gamma, beta);
    check_lsp_stability(&lsp);
  }
  std::vector<double> mgc = lsp2mgc(lsp, alpha, use_log_gain, stage, gamma);
  std::vector<double> cc = gnorm(mc2b(mgc.data(), mgc.size(), alpha), gamma);
  for (size_t i = 1; i < cc.size(); ++i) cc[i] *= gamma;
  return cc;
}

// MelGeneralizedLogSpectrumApproximation::dff (mglsa.rs:21-41)
static void mglsa_dff(std::vector<double>& d, double* x, double alpha, const std::vector<double>& c) {
  const double aa = 1.0 - alpha * alpha;
  double y = d[0] * c[1];
  for (size_t i = 1; i + 1 < c.size(); ++i) {
    d[i] += alpha * (d[i + 1] - d[i - 1]);
    y += d[i] * c[i + 1];
  }
  *x -= y;
  for (size_t i = c.size() - 1; i >= 1; --i) d[i] = d[i - 1];
  d[0] = alpha * d[0] + aa * *x;
}

Vocoder::Vocoder(size_t nmcp_, size_t nlpf_, size_t stage_, bool use_log_gain_, size_t rate_, double alpha_,
                 double beta_, double volume_, size_t fperiod_, bool fma_fir_)
    : nmcp(nmcp_), nlpf(nlpf_), rate(rate_), fperiod(fperiod_), alpha(alpha_), beta(beta_),
      volume(volume_), fma_fir(fma_fir_), stage(stage_), use_log_gain(use_log_gain_) {
  for (auto& v : d21) v.assign(nmcp, 0.0);
  mglsa_d.assign(stage, std::vector<double>(nmcp, 0.0));  // mglsa.rs:9-13
  ring.assign(nlpf, 0.0);
}

void Vocoder::synthesize(double lf0, const double* spectrum, const double* lpf, double* rawdata) {
  // mod.rs:73-77
  double p = 0.0;
  if (lf0 != kNoData) {
    double l = lf0 < kMinLf0 ? kMinLf0 : (lf0 > kMaxLf0 ? kMaxLf0 : lf0);
    p = (double)rate / std::exp(l);
  }
  if (is_first) {  // mod.rs:79-109
    is_first = false;
    coefficients = stage == 0 ? mc2b(spectrum, nmcp, alpha)
                              : lsp_frame_coefficients(spectrum, nmcp, alpha, use_log_gain, stage, beta, false);
  }
  std::vector<double> cc;
  if (stage == 0) {  // Stage::Zero arm, mod.rs:112-141
    std::vector<double> cep(spectrum, spectrum + nmcp);
    postfilter_mcp(&cep, alpha, beta);
    cc = mc2b(cep.data(), nmcp, alpha);
  } else {  // Stage::NonZero arm, mod.rs:142-176
    cc = lsp_frame_coefficients(spectrum, nmcp, alpha, use_log_gain, stage, beta, true);
  }
  std::vector<double> cinc(nmcp);
  for (size_t k = 0; k < nmcp; ++k) cinc[k] = (cc[k] - coefficients[k]) / (double)fperiod;

  // Excitation::start (excitation.rs:25-33)
  if (pitch_of_curr_point != 0.0 && p != 0.0) {
    pitch_inc_per_point = (p - pitch_of_curr_point) / (double)fperiod;
  } else {
    pitch_inc_per_point = 0.0;
    pitch_of_curr_point = p;
    pitch_counter = p;
  }

  for (size_t i = 0; i < fperiod; ++i) {
    // Excitation::get (excitation.rs:67-100)
    double x;
    double tap_pulse = 0.0;
    if (nlpf > 0) {
      double noise = random.nrandom();
      size_t anti = (ring_index + (nlpf - 1) / 2) % nlpf;
      if (pitch_of_curr_point == 0.0) {
        ring[anti] += noise;  // unvoiced_frame
      } else {
        pitch_counter += 1.0;
        double pulse = 0.0;
        if (pitch_counter >= pitch_of_curr_point) {
          pitch_counter -= pitch_of_curr_point;
          pulse = std::sqrt(pitch_of_curr_point);
        }
        tap_pulse = pulse;
        // voiced_frame (excitation.rs:48-64)
        ring[anti] += noise;
        double c = pulse - noise;
        size_t nright = nlpf - ring_index;
        for (size_t k = 0; k < nright; ++k) ring[ring_index + k] += c * lpf[k];
        for (size_t k = 0; k < ring_index; ++k) ring[k] += c * lpf[nright + k];
        pitch_of_curr_point += pitch_inc_per_point;
      }
      x = ring[ring_index];
      ring[ring_index] = 0.0;
      ring_index += 1;
      if (ring_index >= nlpf) ring_index = 0;
    } else if (pitch_of_curr_point == 0.0) {
      x = random.nrandom();
    } else {
      pitch_counter += 1.0;
      if (pitch_counter >= pitch_of_curr_point) {
        pitch_counter -= pitch_of_curr_point;
        x = std::sqrt(pitch_of_curr_point);
        tap_pulse = x;
      } else {
        x = 0.0;
      }
      pitch_of_curr_point += pitch_inc_per_point;
    }
    if (excitation_tap) excitation_tap->push_back(x);

    if (pulse_tap) pulse_tap->push_back(tap_pulse);

    if (stage != 0) {  // mod.rs:161-168
      x *= coefficients[0];
      if (gained_tap) gained_tap->push_back(x);
      for (auto& dd : mglsa_d) mglsa_dff(dd, &x, alpha, coefficients);  // mglsa.rs:15-19
      for (size_t k = 0; k < nmcp; ++k) coefficients[k] += cinc[k];
      rawdata[i] = x * volume;
      continue;
    }
    if (x != 0.0) x *= std::exp(coefficients[0]);  // mod.rs:129-131
    if (gained_tap) gained_tap->push_back(x);

    // df1 (mlsa.rs:54-66)
    {
      const double aa = 1.0 - alpha * alpha;
      double out = 0.0;
      for (size_t k = 5; k >= 1; --k) {
        d11[k] = aa * d12[k - 1] + alpha * d11[k];
        d12[k] = d11[k] * coefficients[1];
        double v = d12[k] * PPADE[k];
        x += (k & 1) ? v : -v;
        out += v;
      }
      d12[0] = x;
      x += out;
    }
    // df2 (mlsa.rs:69-79)
    {
      double out = 0.0;
      for (size_t k = 5; k >= 1; --k) {
        d22[k] = fir(d21[k - 1].data(), nmcp, d22[k - 1], alpha, coefficients.data(), fma_fir);
        double v = d22[k] * PPADE[k];
        x += (k & 1) ? v : -v;
        out += v;
      }
      d22[0] = x;
      x += out;
    }
    for (size_t k = 0; k < nmcp; ++k) coefficients[k] += cinc[k];  // mod.rs:133-135
    rawdata[i] = x * volume;
  }
  pitch_of_curr_point = p;  // Excitation::end
  coefficients = cc;
}

// SpeechGenerator::generate_all (speech.rs:87-96)
std::vector<double> generate_all(Vocoder* v, size_t nframes, const double* mcp, const double* lf0,
                                 const double* lpf) {
  std::vector<double> buf(nframes * v->fperiod, 0.0);
  for (size_t t = 0; t < nframes; ++t)
    v->synthesize(lf0[t], mcp + t * v->nmcp, lpf + t * v->nlpf, buf.data() + t * v->fperiod);
  return buf;
}

}  // namespace oracle
