// Host check of cr_exp (jbonsai_b200/csrc/cuda/cr_exp.h): prints x and cr_exp(x), std::exp(x) as hex bit patterns
// for tools/cr_exp_check.py, which compares them with 80-digit decimal arithmetic.
// Build: g++ -O2 -std=c++17 -ffp-contract=off -mfma -o /tmp/cr_exp_check tools/cr_exp_check.cc
#include <cstdio>
#include <cstdlib>
#include "../jbonsai_b200/csrc/cuda/cr_exp.h"
#define __constant__ static const
#include "../jbonsai_b200/csrc/cuda/exp_table.inc"
int main(int argc, char** argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 100000;
  unsigned long long st = 88172645463325252ull;
  for (int i = 0; i < n; ++i) {
    st ^= st << 13; st ^= st >> 7; st ^= st << 17;
    const double u = (double)(st >> 11) / 9007199254740992.0;
    const double x = 2.995732273553991 + u * (9.903487552536127 - 2.995732273553991);
    const double a = jbk::cr_exp(x, c_exp2_64, kLn2_64_hi, kLn2_64_mid, kLn2_64_lo, k64_Ln2), b = std::exp(x);
    unsigned long long xa, ya, yb;
    memcpy(&xa, &x, 8); memcpy(&ya, &a, 8); memcpy(&yb, &b, 8);
    printf("%llx %llx %llx\n", xa, ya, yb);
  }
}
