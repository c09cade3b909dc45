// The pitch counter of the excitation after m steps of "cnt += 1.0" (src/vocoder/excitation.rs:72-86), without
// walking them. Repeated "+= 1.0" is exact while the sum stays in the binade of the counter (both are multiples of its
// ulp) and rounds only on the step that crosses into the next binade, where the grid gets coarser; below 1 every
// step may round. So the m steps collapse into one exact addition per binade plus the crossing steps themselves,
// each done with the same floating-point addition the reference does.
// Host and device (k_pulse_warp in vocoder.cu; tests/test_cnt_advance.py checks it against the walk on the host).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#ifdef __CUDACC__
#define JB_CA_HD __host__ __device__ __forceinline__
#else
#define JB_CA_HD inline
#endif

namespace jbk {

JB_CA_HD double next_power_of_two_above(double x) {  // 2^(e + 1) for x in [2^e, 2^(e + 1)), x >= 1
  int64_t bits;
  memcpy(&bits, &x, 8);
  bits = (bits & 0x7ff0000000000000ll) + 0x0010000000000000ll;
  double r;
  memcpy(&r, &bits, 8);
  return r;
}

JB_CA_HD double cnt_advance(double cnt, int m) {  // cnt after m times "cnt += 1.0"
  {
    // If cnt + m is exact, every partial sum cnt + j is too (it lies in the same or a lower binade, on a finer
    // grid), so the m single steps round nowhere. After a pulse the counter is a difference of two values near the
    // period, a multiple of the period's ulp: for periods above 256 samples that covers nearly every call.
    const double dm = (double)m, r = cnt + dm, bb = r - cnt;
    if ((cnt - (r - bb)) + (dm - bb) == 0.0) return r;  // TwoSum: the rounding error of cnt + m
  }
  while (m > 0) {
    if (!(cnt >= 1.0)) {  // below 1 the sum may need more bits than it has: a plain step, rounded like the reference
      cnt += 1.0;
      --m;
      continue;
    }
    const double limit = next_power_of_two_above(cnt);
    const double room = limit - cnt;       // exact
    const int kmax = (int)ceil(room) - 1;  // steps that stay below the limit: all exact
    if (kmax >= m) {
      cnt += (double)m;
      m = 0;
    } else {
      cnt += (double)kmax;
      cnt += 1.0;  // the crossing step rounds to the coarser grid, as in the reference
      m -= kmax + 1;
    }
  }
  return cnt;
}

}  // namespace jbk
