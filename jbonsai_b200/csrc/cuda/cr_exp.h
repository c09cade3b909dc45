// exp() rounded correctly (to nearest) for the pitch period, p = rate / exp(lf0) (src/vocoder/mod.rs:73-77).
//
// Why: the north star lists pulse positions under "bit-exact", the period comes from exp(), and CUDA's exp differs
// from glibc's in ~5.7 % of the arguments by one or two ulps (tools/exp_sweep.py, profiles/r2_exp_sweep.json). glibc's
// exp is within 0.51 ulp of the true value, i.e. it IS the correctly rounded result except when the true value sits
// within 0.01 ulp of a rounding boundary; this routine evaluates exp in double-double (error < 2^-75) and rounds once,
// so the two agree wherever glibc rounds correctly. Argument range: |x| < 700 (the caller clamps lf0 to [3.0, 9.9]).
//   x = n ln2/64 + t, |t| <= ln2/128;  exp(x) = 2^(n div 64) * 2^((n mod 64)/64) * exp(t)
// Usable on the host too (tools/cr_exp_check.cc validates it against 80-digit arithmetic).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#ifdef __CUDACC__
#define JB_HD __host__ __device__ __forceinline__
#else
#define JB_HD inline
#endif

namespace jbk {

JB_HD void two_sum(double a, double b, double& s, double& e) {
  s = a + b;
  const double bb = s - a;
  e = (a - (s - bb)) + (b - bb);
}

// `tab` = c_exp2_64 (exp_table.inc): 2^(j/64) as hi, lo pairs; the three parts of ln2/64 come with it.
JB_HD double cr_exp(double x, const double (*tab)[2], double c_hi, double c_mid, double c_lo, double inv_c) {
  const double n = rint(x * inv_c);
  // t = x - n ln2/64 as th + tl: n c_hi is exact (c_hi ends in 12 zero bits), the difference cancels exactly
  const double t_hi = fma(-n, c_hi, x);
  const double m_h = n * c_mid, m_l = fma(n, c_mid, -m_h);
  double th, tl;
  two_sum(t_hi, -m_h, th, tl);
  tl -= m_l + n * c_lo;
  // e = exp(t) - 1 = th + th^2/2 + [tl + th tl + th^3/6 + ...], |t| <= 0.0055: the bracket is < 3e-8
  const double s2 = th * th, s2e = fma(th, th, -s2);
  const double poly = th * s2 * (1.0 / 6 + th * (1.0 / 24 + th * (1.0 / 120 + th * (1.0 / 720 + th * (1.0 / 5040 + th * (1.0 / 40320))))));
  double eh, el;
  two_sum(th, 0.5 * s2, eh, el);
  el += tl + (0.5 * s2e + th * tl) + poly;
  // (Th + Tl) (1 + eh + el), rounded once
  const int ni = (int)n, j = ni & 63, m = (ni - j) / 64;
  const double Th = tab[j][0], Tl = tab[j][1];
  const double p = Th * eh, pe = fma(Th, eh, -p);
  double s, se;
  two_sum(Th, p, s, se);
  const double r = s + (se + pe + Tl + Th * el + Tl * eh);
  // times 2^m: exact for results in the normal range
  int64_t bits;
  memcpy(&bits, &r, 8);
  bits += (int64_t)m << 52;
  double out;
  memcpy(&out, &bits, 8);
  return out;
}

}  // namespace jbk
