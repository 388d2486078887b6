// stk_fastmath.cuh — short-dependency-chain exp / log for LATENCY-bound kernels.
//
// The sequential part of a transition kernel (e.g. models/coal.cu, coal_rj2) runs one warp per scheduler; its time is
// the length of the dependent instruction chain, and CUDA's exp/log (≈ 40-50 instructions, a ~25-deep Horner chain
// with 64-bit immediates) dominate that chain.  These versions trade a 5 KB table in shared memory for depth:
//   fast_log:  renormalise m to [0.75, 1.5), pick one of 192 cells, r = m * (1/c) - 1 in ONE fma (|r| < 2^-8),
//              log m = log c + log1p(r) with a degree-7 polynomial; the two cells next to 1 use c = 1 exactly so
//              that log(x) keeps full RELATIVE accuracy as x -> 1.
//   fast_exp:  x = (256 k + j) ln2/256 + r, |r| < ln2/512, e^x = 2^k * 2^(j/256) * (1 + r + ... + r^5/120).
// Measured against libm (tests/test_fastmath_cpu.py compiles this header for the host): <= 2 ulp over the ranges the
// kernels use; arguments outside the fast range fall back to the library function.
// Plain C++ so that g++ (tests) and NVRTC (generated modules) can both include it.
#ifndef STK_FASTMATH_CUH
#define STK_FASTMATH_CUH

#if defined(__CUDA_ARCH__) || defined(__CUDACC_RTC__)
#define STK_FM __device__ __forceinline__
#define STK_FM_BITS(x) ((unsigned long long)__double_as_longlong(x))
#define STK_FM_DBL(u) __longlong_as_double((long long)(u))
#define STK_FM_FMA(a, b, c) fma(a, b, c)
#else
#include <cmath>
#include <cstring>
#define STK_FM inline
static inline unsigned long long stk_fm_bits(double x) { unsigned long long u; std::memcpy(&u, &x, 8); return u; }
static inline double stk_fm_dbl(unsigned long long u) { double x; std::memcpy(&x, &u, 8); return x; }
#define STK_FM_BITS(x) stk_fm_bits(x)
#define STK_FM_DBL(u) stk_fm_dbl(u)
#define STK_FM_FMA(a, b, c) std::fma(a, b, c)
using std::exp; using std::exp2; using std::log;
#endif

namespace stk {

constexpr int kFmLogCells = 192;                        // cells of width 1/256 over [0.75, 1.5)
constexpr int kFmExpCells = 256;
constexpr int kFmTableDoubles = 2 * kFmLogCells + kFmExpCells;   // (1/c, log c) pairs, then 2^(j/256)

// fill `tab` (kFmTableDoubles doubles, 16-byte aligned); entries idx, idx + stride, ... (call with lane / 32 in a warp)
STK_FM void fm_build_tables(double* tab, int idx, int stride) {
  for (int k = idx; k < kFmLogCells; k += stride) {
    double c = 0.75 + ((double)k + 0.5) / 256.0;
    double inv = 1.0 / c, lc = -log(inv);               // log of the ROUNDED reciprocal's inverse
    if (k == 63 || k == 64) { inv = 1.0; lc = 0.0; }    // [1 - 1/256, 1 + 1/256): r = m - 1 exactly
    tab[2 * k] = inv; tab[2 * k + 1] = lc;
  }
  for (int j = idx; j < kFmExpCells; j += stride) tab[2 * kFmLogCells + j] = exp2((double)j / 256.0);
}

// branch-free core of fast_log for x known to be a positive normal number (no fallback path: the caller patches the
// other cases with a select).  Same arithmetic as fast_log, so the results are bit-identical where both apply.
STK_FM double fast_log_core(double x, const double* tab) {
  const unsigned long long b = STK_FM_BITS(x);
  const int hi = (int)(b >> 32);
  int e = (hi >> 20) - 1023;
  const int up = (hi >> 19) & 1;
  e += up;
  const double m = STK_FM_DBL((b & 0x000fffffffffffffULL) | ((unsigned long long)(1023 - up) << 52));
  const int top = (hi >> 12) & 0xff;
  const int cell = (up ? (top >> 1) - 64 : 64 + top) & 255;          // masked: a wild input cannot index outside the table
  const double inv = tab[2 * cell], lc = tab[2 * cell + 1];
  const double r = STK_FM_FMA(m, inv, -1.0);
  const double r2 = r * r;
  const double p01 = STK_FM_FMA(r, 0.33333333333333333, -0.5);
  const double p23 = STK_FM_FMA(r, 0.2, -0.25);
  const double p45 = STK_FM_FMA(r, 0.14285714285714285, -0.16666666666666666);
  const double r4 = r2 * r2;
  const double q = STK_FM_FMA(r2, p23, p01);
  const double s = STK_FM_FMA(r4, p45, q);
  const double lp = STK_FM_FMA(r2, s, r);
  const double ed = (double)e;
  const double hi_part = STK_FM_FMA(ed, 0.693147180369123816490, lc);
  return hi_part + STK_FM_FMA(ed, 1.90821492927058770002e-10, lp);
}
STK_FM bool fast_log_ok(double x) {          // positive normal finite
  const int hi = (int)(STK_FM_BITS(x) >> 32);
  return hi >= 0x00100000 && hi < 0x7ff00000;
}
STK_FM double fast_log(double x, const double* tab) {
  const unsigned long long b = STK_FM_BITS(x);
  const int hi = (int)(b >> 32);
  if (hi < 0x00100000 || hi >= 0x7ff00000) return log(x);          // zero, subnormal, negative, inf, nan
  int e = (hi >> 20) - 1023;
  const int up = (hi >> 19) & 1;                                     // m >= 1.5 -> halve it
  e += up;
  const double m = STK_FM_DBL((b & 0x000fffffffffffffULL) | ((unsigned long long)(1023 - up) << 52));   // [0.75, 1.5)
  // cell: m in [1, 1.5): 64 + top 7 mantissa bits (top bit is 0);  m in [0.75, 1): top 7 bits of 0.1xxxxxx - 64
  const int top = (hi >> 12) & 0xff;                                 // top 8 mantissa bits
  const int cell = up ? (top >> 1) - 64 : 64 + top;
  const double inv = tab[2 * cell], lc = tab[2 * cell + 1];
  const double r = STK_FM_FMA(m, inv, -1.0);
  // log1p(r) = r - r^2/2 + r^3/3 - ... - r^6/6 + r^7/7, Estrin-style to shorten the chain
  const double r2 = r * r;
  const double p01 = STK_FM_FMA(r, 0.33333333333333333, -0.5);       // -1/2 + r/3
  const double p23 = STK_FM_FMA(r, 0.2, -0.25);                      // -1/4 + r/5
  const double p45 = STK_FM_FMA(r, 0.14285714285714285, -0.16666666666666666);   // -1/6 + r/7
  const double r4 = r2 * r2;
  const double q = STK_FM_FMA(r2, p23, p01);                         // deg <= 3 part of the r^2-scaled series
  const double s = STK_FM_FMA(r4, p45, q);
  const double lp = STK_FM_FMA(r2, s, r);                            // r + r^2 (...)
  const double ed = (double)e;
  const double hi_part = STK_FM_FMA(ed, 0.693147180369123816490, lc);           // ln2 head: 32 significant bits, exact product
  return hi_part + STK_FM_FMA(ed, 1.90821492927058770002e-10, lp);              // ln2 tail
}

STK_FM double fast_exp(double x, const double* tab) {
  if (!(x > -700.0 && x < 700.0)) return exp(x);
  const double magic = 6755399441055744.0;                           // 1.5 * 2^52: round-to-nearest integer in the low bits
  const double t = STK_FM_FMA(x, 369.3299304675746, magic);           // 256 / ln 2
  const int ki = (int)(unsigned)STK_FM_BITS(t);
  const double kd = t - magic;
  double r = STK_FM_FMA(kd, -0.002707606170588406, x);               // ln2/256 head: 30 significant bits, kd * head exact
  r = STK_FM_FMA(kd, -3.4738805620651082e-12, r);                    // ln2/256 tail
  const double tj = tab[2 * kFmLogCells + (ki & 255)];
  const double r2 = r * r;
  const double a = STK_FM_FMA(r, 0.16666666666666666, 0.5);          // 1/2 + r/6
  const double c = STK_FM_FMA(r, 0.0083333333333333332, 0.041666666666666664);   // 1/24 + r/120
  const double p = STK_FM_FMA(r2, STK_FM_FMA(r2, c, a), r);          // r + r^2/2 + ... + r^5/120
  const double v = STK_FM_FMA(tj, p, tj);
  return STK_FM_DBL(STK_FM_BITS(v) + ((unsigned long long)(long long)(ki >> 8) << 52));
}

}  // namespace stk
#endif  // STK_FASTMATH_CUH
