// rng.cuh — per-env Philox4x32-10 request stream (device side).
//
// Replaces the reference's CPython random.Random (MT19937) stream
// (optical_network_env.py:266-271; the five draws of _next_service,
// rmsa_env.py:646-659) in ORLG_SRC_PHILOX mode.  Stream definition
// (DESIGN.md "Philox request source"): request r of global env g uses
//   R0 = philox(ctr=(r.lo, r.hi, g, 0), key=(seed.lo, seed.hi))
//   R1 = philox(ctr=(r.lo, r.hi, g, 1), key)
//   u_arrival = res53(R0.x, R0.y), u_holding = res53(R0.z, R0.w)
//   u_src = R1.x/2^32, u_dst = R1.y/2^32, u_bitrate = R1.z/2^32
// expovariate(l) = -log(1-u)/l with a log made of IEEE add/mul/div only, so
// results do not depend on any libm.  Compile with --fmad=false.
#pragma once
#include <cstdint>

namespace orlg {

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; r++) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

// natural log, fdlibm-style reduction and polynomial, plain IEEE operations
__device__ __forceinline__ double det_log(double x) {
  const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10,
               Lg1 = 6.666666666666735130e-01, Lg2 = 3.999999999940941908e-01,
               Lg3 = 2.857142874366239149e-01, Lg4 = 2.222219843214978396e-01,
               Lg5 = 1.818357216161805012e-01, Lg6 = 1.531383769920937332e-01,
               Lg7 = 1.479819860511658591e-01;
  unsigned long long bits = (unsigned long long)__double_as_longlong(x);
  int k = 0;
  if ((bits >> 52) == 0) {
    x = x * 18014398509481984.0;
    bits = (unsigned long long)__double_as_longlong(x);
    k = -54;
  }
  k += (int)(bits >> 52) - 1023;
  unsigned long long mant = bits & 0x000FFFFFFFFFFFFFull;
  if (mant >= 0x6A09E667F3BCDull) {
    k += 1;
    bits = mant | 0x3FE0000000000000ull;
  } else {
    bits = mant | 0x3FF0000000000000ull;
  }
  double m = __longlong_as_double((long long)bits);
  double f = m - 1.0;
  double s = f / (2.0 + f);
  double z = s * s;
  double w = z * z;
  double t1 = w * (Lg2 + w * (Lg4 + w * Lg6));
  double t2 = z * (Lg1 + w * (Lg3 + w * (Lg5 + w * Lg7)));
  double R = t2 + t1;
  double hfsq = 0.5 * f * f;
  double dk = (double)k;
  return dk * ln2_hi - ((hfsq - (s * (hfsq + R) + dk * ln2_lo)) - f);
}

__device__ __forceinline__ double res53(uint32_t a, uint32_t b) {
  return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) / 9007199254740992.0;
}

// bisect.bisect_right(cum, x, 0, hi) as used by random.choices.  cum is non-decreasing, so the
// insertion point equals the number of entries <= x: hi independent loads instead of a
// dependent binary-search chain.
__device__ __forceinline__ int bisect_right(const double* __restrict__ cum, double x, int hi) {
  int c = 0;
#pragma unroll 8
  for (int i = 0; i < hi; i++) c += (x >= __ldg(cum + i)) ? 1 : 0;
  return c;
}

// The same insertion point from a guide table: the 32-bit draw R selects bucket R >> 26, guide[bucket] is the
// insertion point of the bucket's smallest value (so it is <= the answer for every x of the bucket, x being a
// monotone function of R), and a short forward scan finishes.  One byte load and ~1 compare instead of `hi`.
__device__ __forceinline__ int bisect_right_guided(const double* __restrict__ cum, const unsigned char* __restrict__ guide,
                                                   uint32_t R, double x, int hi) {
  int c = __ldg(guide + (R >> 26));
  while (c < hi && x >= __ldg(cum + c)) c++;
  return c;
}

}  // namespace orlg
