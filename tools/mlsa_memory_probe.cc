// TEST TOOLING (links oracle/liboracle.so): MLSA filter alone, from any frame with a zero state.
// Built and driven by tools/mlsa_memory_probe.py; measures how fast the filter forgets its state.
// MLSA-only filter with selectable start frame and zero state; uses oracle::fir
#include "../oracle/oracle.h"
#include <vector>
#include <cstring>
static const double PP[6] = {1.00000000000, 0.49993910000, 0.11070980000, 0.01369984000, 0.00095648530, 0.00003041721};
extern "C" void mlsa_from(int f0, int f1, int fp, double alpha, const double* mcp /*[T][35]*/, const double* x /*[T*fp] gained*/, double* out /*[ (f1-f0)*fp ]*/) {
  const int nm = 35;
  double d11[6] = {0}, d12[6] = {0}, d22[6] = {0};
  std::vector<double> d21[6];
  for (auto& v : d21) v.assign(nm, 0.0);
  std::vector<double> c = oracle::mc2b(mcp + (size_t)(f0 > 0 ? f0 - 1 : 0) * nm, nm, alpha);
  for (int f = f0; f < f1; ++f) {
    std::vector<double> cc = oracle::mc2b(mcp + (size_t)f * nm, nm, alpha);
    std::vector<double> cinc(nm);
    for (int k = 0; k < nm; ++k) cinc[k] = (cc[k] - c[k]) / (double)fp;
    for (int i = 0; i < fp; ++i) {
      double xx = x[(size_t)f * fp + i];
      { const double aa = 1.0 - alpha * alpha; double o = 0.0;
        for (int k = 5; k >= 1; --k) { d11[k] = aa * d12[k-1] + alpha * d11[k]; d12[k] = d11[k] * c[1]; double v = d12[k] * PP[k]; xx += (k & 1) ? v : -v; o += v; }
        d12[0] = xx; xx += o; }
      { double o = 0.0;
        for (int k = 5; k >= 1; --k) { d22[k] = oracle::fir(d21[k-1].data(), nm, d22[k-1], alpha, c.data(), true); double v = d22[k] * PP[k]; xx += (k & 1) ? v : -v; o += v; }
        d22[0] = xx; xx += o; }
      for (int k = 0; k < nm; ++k) c[k] += cinc[k];
      out[(size_t)(f - f0) * fp + i] = xx;
    }
    c = cc;
  }
}
