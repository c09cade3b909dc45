// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h). Compile with -ffp-contract=off.
// Literal restatement of src/mlpg_adjust/{mod,mask,mlpg}.rs and model/mean_vari.rs.
#include <cmath>

#include "oracle.h"

namespace oracle {

namespace {

struct MeanIvar {
  double mean, ivar;
};

// MeanVari::with_ivar (src/model/mean_vari.rs:21-31)
inline double with_ivar(double vari) {
  if (std::fabs(vari) > 1e19) return 0.0;
  if (std::fabs(vari) < 1e-19) return 1e38;
  return 1.0 / vari;
}

// MlpgMatrix (src/mlpg_adjust/mlpg.rs:14-20)
struct Matrix {
  size_t win_size = 0, length = 0, width = 0;
  std::vector<std::vector<double>> wuw;
  std::vector<double> wum;
};

// MlpgMatrix::calc_wuw_and_wum (mlpg.rs:25-71)
Matrix calc_wuw_and_wum(const std::vector<std::vector<double>>& windows,
                        const std::vector<std::vector<MeanIvar>>& parameters) {
  Matrix m;
  m.length = parameters[0].size();
  size_t maxw = 0;
  for (const auto& w : windows) maxw = w.size() > maxw ? w.size() : maxw;
  m.width = (maxw / 2) * 2 + 1;  // windows.max_width() * 2 + 1
  m.win_size = windows.size();
  m.wuw.assign(m.length, std::vector<double>(m.width, 0.0));
  m.wum.assign(m.length, 0.0);
  const long length = (long)m.length;
  for (long t = 0; t < length; ++t) {
    for (size_t i = 0; i < windows.size(); ++i) {
      const std::vector<double>& win = windows[i];
      const long wwidth = (long)win.size();
      for (long index = wwidth - 1; index >= 0; --index) {  // window.iter_rev(0)
        double coef = win[(size_t)index];
        if (coef == 0.0) continue;
        long idx = t - (index - wwidth / 2);  // t - index.position()
        if (idx < 0 || idx >= length) continue;
        double wu = coef * parameters[i][(size_t)idx].ivar;
        m.wum[(size_t)t] += wu * parameters[i][(size_t)idx].mean;
        for (long inner = wwidth - 1; inner >= index; --inner) {  // window.iter_rev(index)
          double c2 = win[(size_t)inner];
          if (c2 == 0.0) continue;
          long j = inner - index;
          if (t + j >= length) break;
          m.wuw[(size_t)t][(size_t)j] += wu * c2;
        }
      }
    }
  }
  return m;
}

// ldl_factorization (mlpg.rs:79-92)
void ldl_factorization(Matrix* m) {
  auto& wuw = m->wuw;
  for (size_t t = 0; t < m->length; ++t) {
    for (size_t i = 1; i < m->width && i < t + 1; ++i)
      wuw[t][0] -= wuw[t - i][i] * wuw[t - i][i] * wuw[t - i][0];
    for (size_t i = 1; i < m->width; ++i) {
      for (size_t j = 1; j < m->width - i && j < t + 1; ++j)
        wuw[t][i] -= wuw[t - j][j] * wuw[t - j][i + j] * wuw[t - j][0];
      wuw[t][i] /= wuw[t][0];
    }
  }
}

// substitutions (mlpg.rs:95-115)
std::vector<double> substitutions(const Matrix& m) {
  std::vector<double> g(m.length, 0.0);
  for (size_t t = 0; t < m.length; ++t) {
    g[t] = m.wum[t];
    for (size_t i = 1; i < m.width && i < t + 1; ++i) g[t] -= m.wuw[t - i][i] * g[t - i];
  }
  std::vector<double> par(m.length, 0.0);
  for (size_t t = m.length; t-- > 0;) {
    par[t] = g[t] / m.wuw[t][0];
    for (size_t i = 1; i < m.width && i < m.length - t; ++i) par[t] -= m.wuw[t][i] * par[t + i];
  }
  return par;
}

// MlpgGlobalVariance (mlpg.rs:147-292)
struct Gv {
  std::vector<double> par;
  const std::vector<uint8_t>& sw;
  size_t gv_length;
  const Matrix& mtx;

  void calc_gv(double* mean_out, double* vari_out) const {  // :168-188
    double sum = 0.0;
    for (size_t t = 0; t < par.size(); ++t)
      if (sw[t]) sum += par[t];
    double mean = sum / (double)gv_length;
    double acc = 0.0;
    for (size_t t = 0; t < par.size(); ++t)
      if (sw[t]) acc += (par[t] - mean) * (par[t] - mean);
    *mean_out = mean;
    *vari_out = acc / (double)gv_length;
  }
  void conv_gv(double gv_mean) {  // :191-199
    double mean, vari;
    calc_gv(&mean, &vari);
    double ratio = std::sqrt(gv_mean / vari);
    for (size_t t = 0; t < par.size(); ++t)
      if (sw[t]) par[t] = ratio * (par[t] - mean) + mean;
  }
  double calc_hmmobj_derivative(std::vector<double>* g_out) const {  // :200-226
    const size_t length = mtx.length;
    std::vector<double> g(length, 0.0);
    for (size_t t = 0; t < length; ++t) {
      g[t] = mtx.wuw[t][0] * par[t];
      for (size_t i = 1; i < mtx.width; ++i) {
        if (t + i < length) g[t] += mtx.wuw[t][i] * par[t + i];
        if (t + 1 > i) g[t] += mtx.wuw[t - i][i] * par[t - i];
      }
    }
    double w = 1.0 / (double)(mtx.win_size * length);
    double hmmobj = 0.0;
    for (size_t t = 0; t < length; ++t) hmmobj += 1.0 * w * par[t] * (mtx.wum[t] - 0.5 * g[t]);
    *g_out = g;
    return hmmobj;
  }
  void next_step(const std::vector<double>& g, double step, double mean, double vari,
                 double gv_mean, double gv_vari) {  // :227-256
    const size_t length = mtx.length;
    double w = 1.0 / (double)(mtx.win_size * length);
    double dv = -2.0 * gv_vari * (vari - gv_mean) / (double)length;
    for (size_t t = 0; t < length; ++t) {
      double h = -1.0 * w * mtx.wuw[t][0] -
                 1.0 * 2.0 / (double)(length * length) *
                     ((double)(length - 1) * gv_vari * (vari - gv_mean) +
                      2.0 * gv_vari * (par[t] - mean) * (par[t] - mean));
      double next_g;
      if (sw[t])
        next_g = 1.0 / h * (1.0 * w * (-g[t] + mtx.wum[t]) + 1.0 * dv * (par[t] - mean));
      else
        next_g = 1.0 / h * (1.0 * w * (-g[t] + mtx.wum[t]));
      par[t] += step * next_g;
    }
  }
  void parmgen(double gv_mean, double gv_vari) {  // :258-292
    if (gv_length == 0) return;
    double step = 0.1, prev = 0.0;
    conv_gv(gv_mean);
    for (int i = 1; i <= 5; ++i) {
      double mean, vari;
      calc_gv(&mean, &vari);
      double gvobj = -0.5 * 1.0 * vari * gv_vari * (vari - 2.0 * gv_mean);
      std::vector<double> g;
      double hmmobj = calc_hmmobj_derivative(&g);
      double obj = -(hmmobj + gvobj);
      if (i > 1) {
        if (obj > prev)
          step *= 0.5;
        else if (obj < prev)
          step *= 1.2;
      }
      next_step(g, step, mean, vari, gv_mean, gv_vari);
      prev = obj;
    }
  }
};

}  // namespace

// Mask::boundary_distances (mask.rs:51-82)
void boundary_distances(const std::vector<uint8_t>& mask, std::vector<size_t>* left_out,
                        std::vector<size_t>* right_out) {
  size_t n = mask.size();
  left_out->assign(n, 0);
  right_out->assign(n, 0);
  if (n == 0) return;
  size_t left = 0;
  for (size_t f = 0; f < n; ++f) {
    if (mask[f])
      (*left_out)[f] = f - left;
    else
      left = f + 1;
  }
  size_t right = n - 1;
  for (size_t f = n; f-- > 0;) {
    if (mask[f]) {
      (*right_out)[f] = right - f;
    } else {
      if (f == 0) break;
      right = f - 1;
    }
  }
}

// MlpgAdjust::create (mod.rs:51-95) + MlpgMatrix::par (mlpg.rs:118-143)
std::vector<double> mlpg_create(const MlpgInput& in, size_t* nframes) {
  const size_t S = in.nstates, dim = in.dim, nwin = in.windows.size(), L = dim * nwin;
  // Mask::create (mask.rs:20-28)
  std::vector<uint8_t> mask;
  std::vector<size_t> frame_state;
  for (size_t s = 0; s < S; ++s)
    for (int32_t k = 0; k < in.durations[s]; ++k) {
      mask.push_back(in.msd[s] > in.msd_threshold);
      frame_state.push_back(s);
    }
  const size_t T = mask.size();
  *nframes = T;
  std::vector<size_t> left, right;
  boundary_distances(mask, &left, &right);
  std::vector<double> pars(T * dim, 0.0);

  std::vector<uint8_t> gv_switch;  // duration-expanded, mask-filtered (mlpg.rs:129-134)
  if (in.has_gv)
    for (size_t t = 0; t < T; ++t)
      if (mask[t]) gv_switch.push_back(in.gv_switch[frame_state[t]]);

  for (size_t d = 0; d < dim; ++d) {
    std::vector<std::vector<MeanIvar>> parameters(nwin);
    for (size_t w = 0; w < nwin; ++w) {
      const size_t m = dim * w + d;
      const size_t wwidth = in.windows[w].size();
      const size_t lw = wwidth / 2, rw = wwidth - lw - 1;
      for (size_t t = 0; t < T; ++t) {
        if (!mask[t]) continue;
        const size_t s = frame_state[t];
        MeanIvar mi{in.mean[s * L + m], with_ivar(in.vari[s * L + m])};
        bool lb = left[t] < lw, rb = right[t] < rw;
        if ((lb || rb) && w != 0) mi.ivar = 0.0;
        parameters[w].push_back(mi);
      }
    }
    Matrix mtx = calc_wuw_and_wum(in.windows, parameters);
    std::vector<double> par;
    if (in.has_gv) {
      Matrix before = mtx;
      ldl_factorization(&mtx);
      par = substitutions(mtx);
      size_t gvlen = 0;
      for (uint8_t b : gv_switch) gvlen += b ? 1 : 0;
      Gv gv{par, gv_switch, gvlen, before};
      gv.parmgen(in.gv_mean[d] * in.gv_weight, in.gv_vari[d]);
      par = gv.par;
    } else {
      ldl_factorization(&mtx);
      par = substitutions(mtx);
    }
    size_t k = 0;
    for (size_t t = 0; t < T; ++t) pars[t * dim + d] = mask[t] ? par[k++] : kNoData;  // Mask::fill
  }
  return pars;
}

}  // namespace oracle
