// Host check of jbk::cnt_advance (csrc/cuda/cnt_advance.h, used by k_pulse_warp) against the walk it replaces:
// m times "cnt += 1.0" (src/vocoder/excitation.rs:72-86). Prints the number of cases and of mismatches.
// Build: g++ -O2 -std=c++17 -ffp-contract=off -I jbonsai_b200/csrc/cuda -o /tmp/cnt_advance_check tests/native/cnt_advance_check.cc
#include <cstdio>
#include <cstdlib>
#include <random>

#include "cnt_advance.h"

static double walk(double c, int m) {
  for (int i = 0; i < m; ++i) c += 1.0;
  return c;
}

int main(int argc, char** argv) {
  const long n = argc > 1 ? atol(argv[1]) : 2000000;
  std::mt19937_64 rng(12345);
  std::uniform_real_distribution<double> u01(0.0, 1.0);
  long cases = 0, bad = 0;
  auto check = [&](double c, int m) {
    ++cases;
    const double a = jbk::cnt_advance(c, m), b = walk(c, m);
    if (memcmp(&a, &b, 8) != 0) {
      if (++bad <= 5) printf("mismatch: cnt %.17g m %d -> %.17g, walk %.17g\n", c, m, a, b);
    }
  };
  for (long i = 0; i < n; ++i) {
    const int m = (int)(rng() % 301);
    // (a) the counter right after a pulse: (c + 1) - p for a period p and a counter just at it
    const double p = 2.4 + u01(rng) * 2400.0;
    const double c0 = p - u01(rng);  // c0 + 1 >= p
    check((c0 + 1.0) - p, m);
    // (b) any counter below 4096 with a random mantissa
    int64_t bits = (int64_t)(rng() & 0x000fffffffffffffull) | ((int64_t)(1023 - 3 + (int)(rng() % 15)) << 52);
    double x;
    memcpy(&x, &bits, 8);
    check(x, m);
    // (c) just below and at powers of two, integers, tiny values
    const double pw = (double)(1 << (rng() % 12));
    check(std::nextafter(pw, 0.0), m);
    check(pw, m);
    check(pw - 0.5 - u01(rng) * 1e-9, m);
    check(u01(rng) * 1e-300, m);
    check((double)(rng() % 3000), m);
  }
  check(0.0, 0);
  check(0.0, 300);
  check(4.9e-324, 300);
  printf("%ld cases, %ld mismatches\n", cases, bad);
  return bad != 0;
}
