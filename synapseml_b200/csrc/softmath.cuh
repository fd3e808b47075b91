// Deterministic scalar math used by the gradient kernels AND by host code of this library, so that float32 gradients do
// not depend on which libm/libdevice evaluated exp/log2.  Only +,-,*,/ , rint and explicit fma are used; the library is
// compiled with -fmad=false so nothing else is contracted.  (The CPU oracle under oracle/ implements the same published
// algorithm independently; tests assert bit equality.)
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#ifdef __CUDACC__
#define LGB_HD __host__ __device__ __forceinline__
#else
#define LGB_HD inline
#endif

namespace lgb {

LGB_HD double BitsToDouble(uint64_t b) {
#ifdef __CUDA_ARCH__
  return __longlong_as_double(static_cast<long long>(b));
#else
  double d; memcpy(&d, &b, sizeof(d)); return d;
#endif
}
LGB_HD uint64_t DoubleToBits(double d) {
#ifdef __CUDA_ARCH__
  return static_cast<uint64_t>(__double_as_longlong(d));
#else
  uint64_t b; memcpy(&b, &d, sizeof(b)); return b;
#endif
}

// exp(x): k = rint(x*log2e); r = x - k*ln2 (two-term Cody-Waite, fma); degree-13 Taylor by Horner (fma); scale by 2^k.
LGB_HD double SoftExp(double x) {
  if (x > 708.0) x = 708.0;
  if (x < -708.0) x = -708.0;
  const double kf = rint(x * 1.4426950408889634074);
  double r = fma(-kf, 6.93147180369123816490e-01, x);
  r = fma(-kf, 1.90821492927058770002e-10, r);
  double p = 1.0 / 6227020800.0;
  p = fma(p, r, 1.0 / 479001600.0);
  p = fma(p, r, 1.0 / 39916800.0);
  p = fma(p, r, 1.0 / 3628800.0);
  p = fma(p, r, 1.0 / 362880.0);
  p = fma(p, r, 1.0 / 40320.0);
  p = fma(p, r, 1.0 / 5040.0);
  p = fma(p, r, 1.0 / 720.0);
  p = fma(p, r, 1.0 / 120.0);
  p = fma(p, r, 1.0 / 24.0);
  p = fma(p, r, 1.0 / 6.0);
  p = fma(p, r, 0.5);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  const int k = static_cast<int>(kf);
  return p * BitsToDouble(static_cast<uint64_t>(k + 1023) << 52);
}

// log2(x), x > 0 normal: x = m*2^e, m in [sqrt(.5), sqrt(2)); ln m = 2 atanh((m-1)/(m+1)) by its odd series up to s^23.
LGB_HD double SoftLog2(double x) {
  uint64_t bits = DoubleToBits(x);
  int e = static_cast<int>((bits >> 52) & 0x7ff) - 1023;
  double m = BitsToDouble((bits & 0x000fffffffffffffULL) | 0x3ff0000000000000ULL);
  if (m > 1.4142135623730951) { m = m * 0.5; e += 1; }
  const double s = (m - 1.0) / (m + 1.0);
  const double z = s * s;
  double p = 1.0 / 23.0;
  p = fma(p, z, 1.0 / 21.0);
  p = fma(p, z, 1.0 / 19.0);
  p = fma(p, z, 1.0 / 17.0);
  p = fma(p, z, 1.0 / 15.0);
  p = fma(p, z, 1.0 / 13.0);
  p = fma(p, z, 1.0 / 11.0);
  p = fma(p, z, 1.0 / 9.0);
  p = fma(p, z, 1.0 / 7.0);
  p = fma(p, z, 1.0 / 5.0);
  p = fma(p, z, 1.0 / 3.0);
  p = fma(p, z, 1.0);
  const double ln_m = 2.0 * s * p;
  return fma(ln_m, 1.4426950408889634074, static_cast<double>(e));
}

// Counter-based uniform in [0,1) for stochastic rounding of quantized gradients, keyed by (seed, iteration, global row, which).
LGB_HD float HashUniform(uint32_t seed, uint32_t iter, uint32_t row, uint32_t which) {
  uint32_t h = seed * 0x9E3779B1u + iter * 0x85EBCA77u + row * 0xC2B2AE3Du + which * 0x27D4EB2Fu + 0x165667B1u;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; h *= 0x297A2D39u; h ^= h >> 15;
  return static_cast<float>(h >> 8) * (1.0f / 16777216.0f);
}

}  // namespace lgb
