// pow_glibc.cuh — pow(double,double) with the exact operation sequence of glibc 2.39's
// __pow_fma (x86-64 FMA variant, the one CPython's `float ** float` reaches on any host with
// FMA+AVX2 — every B200 host CPU).  Reference call sites: agilerl/components/
// replay_buffer.py:322 (priority ** alpha), :401/:406 ((p * size) ** -beta).
//
// Algorithm: ARM Optimized Routines math/pow.c (glibc sysdeps/ieee754/dbl-64/e_pow.c):
//   log(x) = k ln2 + log(c) + log1p(z/c - 1) in double-double (hi, lo),
//   (ehi, elo) = y * (hi, lo),  result = 2^(k/128) table * exp polynomial.
// Which products and sums are fused (fma) below follows the instruction stream of __pow_fma
// as compiled into the image's libm.so.6 (GCC contracts more than the C source spells out);
// that is what makes the result bit-identical rather than merely < 1 ULP.  Verified on the
// host against libm for 10^8+ inputs by tools/check_pow_host.cpp (same source, fma() is exact
// on both sides), and on the device by tests/test_kernels_gpu.py::test_pow_is_cpython_pow.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define XP_HD __device__ __forceinline__
#define XP_POW_TABLE static __device__ const
#else
#define XP_HD static inline
#define XP_POW_TABLE static const
#endif
#include "pow_tables.cuh"

namespace xp {

#if defined(__CUDACC__)
#define XP_BITS2D(u) __longlong_as_double((long long)(u))
#define XP_D2BITS(d) ((uint64_t)__double_as_longlong(d))
#else
static inline double bits2d(uint64_t u) { double d; __builtin_memcpy(&d, &u, 8); return d; }
static inline uint64_t d2bits(double d) { uint64_t u; __builtin_memcpy(&u, &d, 8); return u; }
#define XP_BITS2D(u) xp::bits2d(u)
#define XP_D2BITS(d) xp::d2bits(d)
#endif
#define XP_TBL(t, i) XP_BITS2D(t[i])
#define XP_TBLU(t, i) (t[i])

// 0: not an integer, 1: odd integer, 2: even integer (e_pow.c checkint)
XP_HD int pow_checkint(uint64_t iy) {
  const int e = (int)((iy >> 52) & 0x7ff);
  if (e < 0x3ff) return 0;
  if (e > 0x3ff + 52) return 2;
  if (iy & ((1ULL << (0x3ff + 52 - e)) - 1)) return 0;
  if (iy & (1ULL << (0x3ff + 52 - e))) return 1;
  return 2;
}

XP_HD bool pow_zeroinfnan(uint64_t i) { return 2 * i - 1 >= 2 * 0x7ff0000000000000ULL - 1; }

// exp(x + xtail) * (-1)^sign with the result scaled by 2^k, k possibly out of range
XP_HD double pow_exp_special(double tmp, uint64_t sbits, uint64_t ki) {
  double scale, y;
  if ((ki & 0x80000000ULL) == 0) {
    sbits -= 1009ULL << 52;
    scale = XP_BITS2D(sbits);
    y = 0x1p1009 * fma(scale, tmp, scale);
    return y;
  }
  sbits += 1022ULL << 52;
  scale = XP_BITS2D(sbits);
  const double st = scale * tmp;  // rounded product, reused (not fused in __pow_fma)
  y = scale + st;
  if (fabs(y) < 1.0) {
    double one = 1.0;
    if (y < 0.0) one = -1.0;
    double lo = (scale - y) + st;
    const double hi = one + y;
    lo = ((one - hi) + y) + lo;
    y = (lo + hi) - one;
    if (y == 0.0) y = XP_BITS2D(sbits & 0x8000000000000000ULL);
  }
  return 0x1p-1022 * y;
}

XP_HD double xp_pow_glibc(double x, double y) {
  uint64_t sign_bias = 0;
  uint64_t ix = XP_D2BITS(x);
  const uint64_t iy = XP_D2BITS(y);
  uint32_t topx = (uint32_t)(ix >> 52);
  const uint32_t topy = (uint32_t)(iy >> 52);
  if (topx - 0x001u >= 0x7ffu - 0x001u || (topy & 0x7ff) - 0x3beu >= 0x43eu - 0x3beu) {
    if (pow_zeroinfnan(iy)) {
      if (2 * iy == 0) return 1.0;
      if (ix == 0x3ff0000000000000ULL) return 1.0;
      if (2 * ix > 2 * 0x7ff0000000000000ULL || 2 * iy > 2 * 0x7ff0000000000000ULL) return x + y;
      if (2 * ix == 2 * 0x3ff0000000000000ULL) return 1.0;
      if ((2 * ix < 2 * 0x3ff0000000000000ULL) == !(iy >> 63)) return 0.0;
      return y * y;
    }
    if (pow_zeroinfnan(ix)) {
      double x2 = x * x;
      if ((ix >> 63) && pow_checkint(iy) == 1) x2 = -x2;
      return (iy >> 63) ? 1.0 / x2 : x2;
    }
    if (ix >> 63) {
      const int yint = pow_checkint(iy);
      if (yint == 0) return XP_BITS2D(0x7ff8000000000000ULL);  // invalid: NaN
      if (yint == 1) sign_bias = 0x800ULL << 7;
      ix &= 0x7fffffffffffffffULL;
      topx &= 0x7ff;
    }
    if ((topy & 0x7ff) - 0x3beu >= 0x43eu - 0x3beu) {
      if (ix == 0x3ff0000000000000ULL) return 1.0;
      if ((topy & 0x7ff) < 0x3beu) return ix > 0x3ff0000000000000ULL ? 1.0 + y : 1.0 - y;
      return ((ix > 0x3ff0000000000000ULL) == (topy < 0x800u)) ? XP_BITS2D(0x7ff0000000000000ULL)
                                                              : 0.0;
    }
    if (topx == 0) {
      ix = XP_D2BITS(x * 0x1p52);
      ix &= 0x7fffffffffffffffULL;
      ix -= 52ULL << 52;
    }
  }

  // ---- log_inline: hi + lo = log(x) --------------------------------------------------
  const uint64_t tmpi = ix - 0x3fe6955500000000ULL;
  const int i = (int)((tmpi >> (52 - 7)) & 127);
  const int k = (int)((int64_t)tmpi >> 52);
  const uint64_t iz = ix - (tmpi & (0xfffULL << 52));
  const double z = XP_BITS2D(iz);
  const double kd = (double)k;
  const double invc = XP_TBL(kPowInvc, i), logc = XP_TBL(kPowLogc, i);
  const double logctail = XP_TBL(kPowLogctail, i);
  const double Ln2hi = XP_TBL(kPowLn2, 0), Ln2lo = XP_TBL(kPowLn2, 1);
  const double A0 = XP_TBL(kPowLogPoly, 0), A1 = XP_TBL(kPowLogPoly, 1);
  const double A2 = XP_TBL(kPowLogPoly, 2), A3 = XP_TBL(kPowLogPoly, 3);
  const double A4 = XP_TBL(kPowLogPoly, 4), A5 = XP_TBL(kPowLogPoly, 5);
  const double A6 = XP_TBL(kPowLogPoly, 6);

  const double t1 = fma(kd, Ln2hi, logc);
  const double lo1 = fma(kd, Ln2lo, logctail);
  const double r = fma(z, invc, -1.0);
  const double ar = r * A0;
  const double p12 = fma(r, A2, A1);
  const double p34 = fma(r, A4, A3);
  const double t2 = r + t1;
  const double lo2 = (t1 - t2) + r;
  const double ar2 = r * ar;
  const double ar3 = r * ar2;
  const double lo3 = fma(ar, r, -ar2);
  const double hi = t2 + ar2;
  const double p56 = fma(r, A6, A5);
  const double lo4 = (t2 - hi) + ar2;
  const double q = fma(ar2, fma(p56, ar2, p34), p12);
  const double lo = fma(ar3, q, ((lo1 + lo2) + lo3) + lo4);
  const double lhi = hi + lo;
  const double llo = (hi - lhi) + lo;

  // ---- y * log(x) in double-double ---------------------------------------------------------
  const double ehi = y * lhi;
  const double elo = fma(y, llo, fma(lhi, y, -ehi));

  // ---- exp_inline(ehi, elo, sign_bias) -------------------------------------------------------
  uint32_t abstop = (uint32_t)(XP_D2BITS(ehi) >> 52) & 0x7ff;
  if (abstop - 0x3c9u >= 0x408u - 0x3c9u) {
    if (abstop - 0x3c9u >= 0x80000000u) {
      const double one = 1.0 + ehi;
      return sign_bias ? -one : one;
    }
    if (abstop >= 0x409u) {
      if (XP_D2BITS(ehi) >> 63) return sign_bias ? -0.0 : 0.0;
      const double inf = XP_BITS2D(0x7ff0000000000000ULL);
      return sign_bias ? -inf : inf;
    }
    abstop = 0;
  }
  const double InvLn2N = XP_TBL(kExpConst, 0), Shift = XP_TBL(kExpConst, 1);
  const double NegLn2hiN = XP_TBL(kExpConst, 2), NegLn2loN = XP_TBL(kExpConst, 3);
  const double C2 = XP_TBL(kExpConst, 4), C3 = XP_TBL(kExpConst, 5);
  const double C4 = XP_TBL(kExpConst, 6), C5 = XP_TBL(kExpConst, 7);
  double kd2 = fma(ehi, InvLn2N, Shift);
  const uint64_t ki = XP_D2BITS(kd2);
  kd2 = kd2 - Shift;
  double rr = fma(kd2, NegLn2loN, fma(kd2, NegLn2hiN, ehi));
  rr = elo + rr;
  const unsigned idx = 2 * (unsigned)(ki % 128);
  const uint64_t top = (ki + sign_bias) << (52 - 7);
  const double tail = XP_BITS2D(XP_TBLU(kExpTab, idx));
  const uint64_t sbits = XP_TBLU(kExpTab, idx + 1) + top;
  const double r2 = rr * rr;
  const double e23 = fma(rr, C3, C2);
  const double e45 = fma(rr, C5, C4);
  const double tmp = fma(e45, r2 * r2, fma(e23, r2, rr + tail));
  if (abstop == 0) return pow_exp_special(tmp, sbits, ki);
  const double scale = XP_BITS2D(sbits);
  return fma(tmp, scale, scale);
}

}  // namespace xp
