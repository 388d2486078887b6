// stk_device.cuh — device library shared by every model module (hand-tuned
// builtins and NVRTC-compiled stkc output): counter-based Philox4x32-10 RNG,
// samplers, closed-form log-densities, warp/sub-warp reductions.
//
// The log-density formulas restate the interpreter's primitive table
//   stochaskell/src/Data/Expression/Const.hs:90-106 and
//   stochaskell/src/Data/Random/Distribution/Abstract.hs:68,77,90,113-115,121,152,161-167
// in log space (NOT log(boost::math::pdf(..)) as CC.hs:206-228 does, which
// underflows to -inf in the tails; see DESIGN.md "parity notes").
//
// NVRTC-safe: no host headers.
#ifndef STK_DEVICE_CUH
#define STK_DEVICE_CUH

#include "stk_abi.h"

namespace stk {

typedef unsigned int u32;
typedef unsigned long long u64;

#define STK_DEV __device__ __forceinline__

constexpr double kLog2Pi = 1.8378770664093454835606594728112;   // log(2*pi)
constexpr double kHalfLog2Pi = 0.91893853320467274178032973640562;
constexpr double kLogPi = 1.1447298858494001741434273513531;
constexpr double kTwoPi = 6.283185307179586476925286766559;

STK_DEV double neg_inf() { return __longlong_as_double(0xfff0000000000000ULL); }
STK_DEV double pos_inf() { return __longlong_as_double(0x7ff0000000000000ULL); }

// ---------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11).  Counter layout used everywhere in this
// backend (and restated bit-for-bit by oracle/stk_oracle.cpp):
//   key = (seed lo32, seed hi32)
//   ctr = (global chain index, transition lo32, transition hi32, slot)
// `slot` counts the 128-bit blocks a chain consumes inside one transition.
// ---------------------------------------------------------------------------
struct u32x4 { u32 x, y, z, w; };

STK_DEV u32x4 philox4x32_10(u32 c0, u32 c1, u32 c2, u32 c3, u32 k0, u32 k1) {
  const u32 M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  const u32 W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    u32 hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    u32 hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    u32 n0 = hi1 ^ c1 ^ k0;
    u32 n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  u32x4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
  return o;
}

// 53-bit uniform in the OPEN interval (0,1): ((hi>>5)*2^26 + (lo>>6) + 0.5) * 2^-53
STK_DEV double u01_from_bits(u32 hi, u32 lo) {
  u64 m = ((u64)(hi >> 5) << 26) | (u64)(lo >> 6);
  return ((double)m + 0.5) * 1.1102230246251565404236316680908e-16;
}

struct Rng {
  u32 k0, k1, c0, c1, c2, slot;
  STK_DEV void init(u64 seed, u64 global_chain, u64 step) {
    k0 = (u32)seed; k1 = (u32)(seed >> 32);
    c0 = (u32)global_chain; c1 = (u32)step; c2 = (u32)(step >> 32);
    slot = 0;
  }
  STK_DEV u32x4 next() { return philox4x32_10(c0, c1, c2, slot++, k0, k1); }
  // one block -> one uniform (words x,y)
  STK_DEV double u01() { u32x4 b = next(); return u01_from_bits(b.x, b.y); }
  // one block -> two uniforms
  STK_DEV void u01x2(double& a, double& b) {
    u32x4 r = next();
    a = u01_from_bits(r.x, r.y); b = u01_from_bits(r.z, r.w);
  }
  // one block -> two standard normals (Box-Muller)
  STK_DEV void normal2(double& a, double& b) {
    double u1, u2; u01x2(u1, u2);
    double r = sqrt(-2.0 * log(u1));
    double s, c; sincospi(2.0 * u2, &s, &c);
    a = r * c; b = r * s;
  }
  STK_DEV double normal() { double a, b; normal2(a, b); return a; }
};

// ---- samplers (each documents how many blocks it consumes) -----------------
// uniform(a,b): 1 block.  CC.hs:280 (uniform_real_distribution)
STK_DEV double draw_uniform(Rng& g, double a, double b) { return a + (b - a) * g.u01(); }
// normal(m,s): 1 block.  CC.hs:278
STK_DEV double draw_normal(Rng& g, double m, double s) { return m + s * g.normal(); }
// discreteUniform(a,b) inclusive: 1 block, exact integer arithmetic.  CC.hs:277
STK_DEV int draw_discrete_uniform(Rng& g, int a, int b) {
  int n = b - a + 1;
  int k = (int)floor(g.u01() * (double)n);
  if (k >= n) k = n - 1;
  return a + k;
}
// bernoulli(p): 1 block.  CC.hs:290-291
STK_DEV int draw_bernoulli(Rng& g, double p) { return g.u01() < p ? 1 : 0; }
// pmf(ps) -> 1-based index: 1 block; linear scan of the running sum, ps need
// not be normalised (std::discrete_distribution normalises, CC.hs:292-293).
STK_DEV int draw_pmf(Rng& g, const double* ps, int n) {
  double tot = 0.0;
  for (int i = 0; i < n; ++i) tot += ps[i];
  double u = g.u01() * tot, acc = 0.0;
  int k = n;
  for (int i = 0; i < n; ++i) { acc += ps[i]; if (u < acc) { k = i + 1; break; } }
  return k;
}
// poisson(l): 1 block when l < 30 (inverse CDF by sequential search, exact
// integer result), else PTRS-free normal approximation is NOT used: we loop
// the product method in blocks so the result stays an exact function of the
// stream.  CC.hs:279
STK_DEV int draw_poisson(Rng& g, double l) {
  if (l < 30.0) {
    double u = g.u01();
    double p = exp(-l), c = p;
    int k = 0;
    while (u >= c && k < 1000) { ++k; p *= l / (double)k; c += p; }
    return k;
  }
  // split: Poisson(l) = sum of m iid Poisson(l/m) with l/m < 30
  int m = (int)(l / 25.0) + 1;
  double lm = l / (double)m;
  int tot = 0;
  for (int j = 0; j < m; ++j) {
    double u = g.u01();
    double p = exp(-lm), c = p;
    int k = 0;
    while (u >= c && k < 1000) { ++k; p *= lm / (double)k; c += p; }
    tot += k;
  }
  return tot;
}
// gamma(shape a, rate b): Marsaglia-Tsang squeeze; consumes 1 block per
// attempt (two normals unused half) + 1 block for the a<1 boost.  CC.hs:294-296
STK_DEV double draw_gamma(Rng& g, double a, double b) {
  double boost = 1.0;
  if (a < 1.0) { boost = pow(g.u01(), 1.0 / a); a += 1.0; }
  double d = a - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
  for (int it = 0; it < 64; ++it) {
    u32x4 r = g.next();
    double u1 = u01_from_bits(r.x, r.y), u2 = u01_from_bits(r.z, r.w);
    // normal from (u1) via inverse-free polar trick is avoided: use Box-Muller
    // cosine branch on (u1,u2) then a THIRD uniform from the next block only
    // when the squeeze fails.
    double rad = sqrt(-2.0 * log(u1));
    double s, cs; sincospi(2.0 * u2, &s, &cs);
    double x = rad * cs;
    double v = 1.0 + c * x;
    if (v <= 0.0) continue;
    v = v * v * v;
    double u = g.u01();
    if (log(u) < 0.5 * x * x + d - d * v + d * log(v)) return boost * d * v / b;
  }
  return boost * d / b;
}
// cauchy(m,s): ratio of two normals, Abstract.hs:58-64.  1 block.
STK_DEV double draw_cauchy(Rng& g, double m, double s) {
  double x, y; g.normal2(x, y); return m + s * (x / y);
}

// inv_gamma(a, b) = 1 / gamma(a, b).  Abstract.hs:72-75
STK_DEV double draw_inv_gamma(Rng& g, double a, double b) { return 1.0 / draw_gamma(g, a, b); }
// geometric(p): failures before the first success.  The reference flips bernoulli(p) coins one by one
// (Abstract.hs:80-85); the same law by inversion of one uniform, k = floor(log u / log(1 - p)): 1 block.
STK_DEV int draw_geometric(Rng& g, double p) {
  const double u = g.u01();
  if (p >= 1.0) return 0;
  const double k = floor(log(u) / log1p(-p));
  return k < 2147483647.0 ? (int)k : 2147483647;
}
// logistic(m, s) = m + s log(u / (1 - u)).  Abstract.hs:101-104.  1 block.
STK_DEV double draw_logistic(Rng& g, double m, double s) { const double u = g.u01(); return m + s * log(u / (1.0 - u)); }
// neg_binomial(a, b): l ~ gamma(a, b), then poisson(l).  Abstract.hs:108-111
STK_DEV int draw_neg_binomial(Rng& g, double a, double b) { const double l = draw_gamma(g, a, b); return draw_poisson(g, l); }
// beta(a, b) = X / (X + Y), X ~ gamma(a, 1), Y ~ gamma(b, 1) (the construction random-fu's Beta uses, Abstract.hs:40-41)
STK_DEV double draw_beta(Rng& g, double a, double b) {
  const double X = draw_gamma(g, a, 1.0), Y = draw_gamma(g, b, 1.0);
  return X / (X + Y);
}
// lkj_corr(v) for a 2 x 2 correlation matrix: the off-diagonal r = 2 beta(v, v) - 1.  Abstract.hs:93-97
STK_DEV double draw_lkj_corr2(Rng& g, double v) { return 2.0 * draw_beta(g, v, v) - 1.0; }

// ---- log densities ---------------------------------------------------------
STK_DEV double log_factorial(int k) { return lgamma(1.0 + (double)k); }  // CC.hs:167-168

STK_DEV double normal_lpdf(double x, double m, double s) {           // Abstract.hs:121
  double z = (x - m) / s;
  return -kHalfLog2Pi - log(s) - 0.5 * z * z;
}
STK_DEV double gamma_lpdf(double x, double a, double b) {            // Abstract.hs:68
  return a * log(b) + (a - 1.0) * log(x) - b * x - lgamma(a);
}
STK_DEV double inv_gamma_lpdf(double x, double a, double b) {        // Abstract.hs:77
  return a * log(b) + (-a - 1.0) * log(x) - b / x - lgamma(a);
}
STK_DEV double geometric_lpdf(int k, double p) {                     // Abstract.hs:90
  return (double)k * log(1.0 - p) + log(p);
}
STK_DEV double neg_binomial_lpdf(int k, double a, double b) {        // Abstract.hs:113-115
  return lgamma(a + (double)k) - lgamma(a) - log_factorial(k) +
         (double)k * log(1.0 / (b + 1.0)) + a * log(b / (b + 1.0));
}
STK_DEV double poisson_lpdf(int k, double l) {                       // Abstract.hs:152
  return (double)k * log(l) - log_factorial(k) - l;
}
STK_DEV double uniform_lpdf(double x, double a, double b) {          // Abstract.hs:161-162
  return (x < a || b < x) ? neg_inf() : log(1.0 / (b - a));
}
STK_DEV double uniform_cdf(double x, double a, double b) {           // Abstract.hs:163-165
  return x < a ? 0.0 : (x > b ? 1.0 : (x - a) / (b - a));
}
STK_DEV double discrete_uniform_lpdf(int x, int a, int b) {          // Abstract.hs:166-167
  return (x < a || b < x) ? neg_inf() : log(1.0 / (double)(b - a + 1));
}
STK_DEV double bernoulli_lpdf(int b, double p) {                     // Const.hs:91
  return log(b ? p : 1.0 - p);
}
STK_DEV double bernoulli_logit_lpdf(int b, double l) {               // Const.hs:92-93, CC.hs:198-200
  double p = 1.0 / (1.0 + exp(-l));
  return log(b ? p : 1.0 - p);
}
STK_DEV double pmf_lpdf(int k, const double* ps) { return log(ps[k - 1]); }  // Const.hs:94
STK_DEV double cauchy_lpdf(double x, double m, double s) {           // stan_math cauchy_lpdf.hpp
  double z = (x - m) / s;
  return -kLogPi - log(s) - log1p(z * z);
}
STK_DEV double logistic_lpdf(double x, double m, double s) {
  double z = (x - m) / s;
  return -z - log(s) - 2.0 * log1p(exp(-z));
}
STK_DEV double beta_lpdf(double x, double a, double b) {
  return lgamma(a + b) - lgamma(a) - lgamma(b) + (a - 1.0) * log(x) + (b - 1.0) * log1p(-x);
}
// 1-based sorted insert position: first j with e < v[j], else n+1.
// CC.hs:192-197 (findSortedInsertIndex)
STK_DEV int find_sorted_insert_index(double e, const double* v, int n) {
  int j = n + 1;
  for (int i = 1; i <= n; ++i) if (e < v[i - 1]) { j = i; break; }
  return j;
}

// ---- vectors of the generated (stkc) kernels ---------------------------------
// A view: 1-based `at`, element stride in elements (n_chains for a state field stored [len][chain], 1 for
// a per-thread buffer).  Matrix<.., Dynamic, 1> of the C++ backend (CC.hs:77), without ownership.
// `at` is bounds-checked: the reference's interpreter evaluates lazily, so a printed IR may index a vector at a position that
// exists only under the guard that uses the value (e.g. `log input!v` with v = -1 "no differing element" in the coal kernel,
// rj-coal.ipynb cell 8 v_0_172..v_0_176); an out-of-range read yields NaN (reals) / 0 (integers) instead of touching memory.
template <class T> STK_DEV T oob_value();
template <> STK_DEV double oob_value<double>() { return __longlong_as_double(0x7ff8000000000000ULL); }
template <> STK_DEV int oob_value<int>() { return 0; }
template <class T>
struct Vec {
  T* p; long long stride; int len;
  STK_DEV Vec() : p(nullptr), stride(1), len(0) {}
  STK_DEV Vec(T* p_, long long s_, int n_) : p(p_), stride(s_), len(n_) {}
  STK_DEV T at(int i) const { return (i >= 1 && i <= len) ? p[(long long)(i - 1) * stride] : oob_value<T>(); }
};
typedef Vec<double> VecD;
typedef Vec<int> VecI;
// length of a vector about to be written into storage of `cap` elements: an overflow is clamped AND reported through the
// module's `err` work field (the runtime turns it into STK_ERR_ARG at the next synchronising call), never silent
STK_DEV int cap_len(int n, int cap, int* err, int code = 1) {
  if (n > cap) { if (err) *err = code; return cap; }
  return n < 0 ? 0 : n;
}
// next-transition feed: make the state vector `dst` hold `src` (no copy when it already does)
template <class T>
STK_DEV void assign_vec(Vec<T>& dst, const Vec<T>& src, int cap, int* err = nullptr) {
  const int n = cap_len(src.len, cap, err, 2);
  if (src.p != dst.p) for (int i = 0; i < n; ++i) dst.p[(long long)i * dst.stride] = src.p[(long long)i * src.stride];
  dst.len = n;
}
template <class T>
STK_DEV void store_vec(const Vec<T>& v, T* dst, long long stride, int cap, int* len_out, int* err = nullptr) {
  const int n = cap_len(v.len, cap, err, 2);
  if (v.p != dst) for (int i = 0; i < n; ++i) dst[(long long)i * stride] = v.p[(long long)i * v.stride];
  *len_out = n;
}
// vector equality (`a == b` on two vectors: same length, same elements)
template <class T>
STK_DEV bool vec_eq(const Vec<T>& a, const Vec<T>& b) {
  if (a.len != b.len) return false;
  for (int i = 1; i <= a.len; ++i) if (!(a.at(i) == b.at(i))) return false;
  return true;
}
// log |det A| of a dense n x n matrix stored row-major with pitch n, by LU with partial pivoting, in place (the interpreter's
// logDet = fst . snd . invlndet, Data/Array/Abstract.hs:329-331; LOG_DET of CC.hs:420-421 covers the positive-definite case).
// A singular matrix gives -inf.
// NOT inlined, single exit: inlined four times into the divergent branches of a generated reversible-jump kernel (one LU per
// move kind, early returns out of the pivot loop), ptxas -O1..-O3 (CUDA 12.9, sm_100a) produced code in which warps whose lanes
// took different branches never reconverged — the kernel hung on B200 while -O0 / -G builds and a host build of the same source
// ran to completion (profiles/r02_coal_ir_hang.md).  As a real function call the control flow of the factorisation is isolated.
__device__ __noinline__ double log_abs_det(double* A, int n) {
  double acc = 0.0;
  bool done = false;
  for (int k = 0; k < n && !done; ++k) {
    int piv = k; double best = fabs(A[k * n + k]);
    for (int i = k + 1; i < n; ++i) { const double v = fabs(A[i * n + k]); if (v > best) { best = v; piv = i; } }
    if (!(best > 0.0)) { acc = best == 0.0 ? neg_inf() : oob_value<double>(); done = true; }
    else {
      if (piv != k) for (int j = k; j < n; ++j) { const double t = A[k * n + j]; A[k * n + j] = A[piv * n + j]; A[piv * n + j] = t; }
      const double d = A[k * n + k];
      acc += log(fabs(d));
      for (int i = k + 1; i < n; ++i) {
        const double f = A[i * n + k] / d;
        if (f != 0.0) for (int j = k + 1; j < n; ++j) A[i * n + j] -= f * A[k * n + j];
      }
    }
  }
  return acc;
}
// ---- small dense linear algebra on per-thread storage (the Eigen one-liners of ccNode, CC.hs:157-166, 221-225; the interpreter's
// hmatrix primitives, Data/Array/Abstract.hs:310-333).  Row-major, pitch = n.  Real function calls (see log_abs_det). -----------
// lower Cholesky factor in place (`chol`, CC.hs:161-162: m.llt().matrixL()); the strict upper triangle is zeroed; a non-positive
// pivot poisons the factor with NaN (the reference's `chol` raises an error there)
__device__ __noinline__ void chol_lower(double* A, int n) {
  bool bad = false;
  for (int j = 0; j < n; ++j) {
    double d = A[j * n + j];
    for (int k = 0; k < j; ++k) d -= A[j * n + k] * A[j * n + k];
    if (!(d > 0.0)) bad = true;
    const double l = sqrt(d);
    A[j * n + j] = l;
    for (int i = j + 1; i < n; ++i) {
      double v = A[i * n + j];
      for (int k = 0; k < j; ++k) v -= A[i * n + k] * A[j * n + k];
      A[i * n + j] = v / l;
    }
    for (int i = 0; i < j; ++i) A[i * n + j] = 0.0;
  }
  if (bad) for (int i = 0; i < n * n; ++i) A[i] = oob_value<double>();
}
// forward substitution L x = b, in place on b (b has m columns, row-major n x m)
__device__ __noinline__ void tri_solve_lower(const double* L, int n, double* b, int m) {
  for (int c = 0; c < m; ++c)
    for (int i = 0; i < n; ++i) {
      double v = b[i * m + c];
      for (int k = 0; k < i; ++k) v -= L[i * n + k] * b[k * m + c];
      b[i * m + c] = v / L[i * n + i];
    }
}
// A X = B by LU with partial pivoting (`<\>`, CC.hs:159-160: colPivHouseholderQr().solve; `inv`, :163-164); A is destroyed, the
// solution replaces B (n x m, row-major).  A singular matrix gives NaN.
__device__ __noinline__ void lu_solve(double* A, int n, double* B, int m) {
  bool bad = false;
  for (int k = 0; k < n; ++k) {
    int piv = k; double best = fabs(A[k * n + k]);
    for (int i = k + 1; i < n; ++i) { const double v = fabs(A[i * n + k]); if (v > best) { best = v; piv = i; } }
    if (!(best > 0.0)) bad = true;
    if (piv != k) {
      for (int j = 0; j < n; ++j) { const double t = A[k * n + j]; A[k * n + j] = A[piv * n + j]; A[piv * n + j] = t; }
      for (int j = 0; j < m; ++j) { const double t = B[k * m + j]; B[k * m + j] = B[piv * m + j]; B[piv * m + j] = t; }
    }
    const double d = A[k * n + k];
    for (int i = k + 1; i < n; ++i) {
      const double f = A[i * n + k] / d;
      if (f != 0.0) {
        for (int j = k + 1; j < n; ++j) A[i * n + j] -= f * A[k * n + j];
        for (int j = 0; j < m; ++j) B[i * m + j] -= f * B[k * m + j];
      }
    }
  }
  for (int c = 0; c < m; ++c)
    for (int i = n - 1; i >= 0; --i) {
      double v = B[i * m + c];
      for (int k = i + 1; k < n; ++k) v -= A[i * n + k] * B[k * m + c];
      B[i * m + c] = v / A[i * n + i];
    }
  if (bad) for (int i = 0; i < n * m; ++i) B[i] = oob_value<double>();
}
// determinant by the same factorisation (`det`)
__device__ __noinline__ double det_lu(double* A, int n) {
  double det = 1.0;
  for (int k = 0; k < n; ++k) {
    int piv = k; double best = fabs(A[k * n + k]);
    for (int i = k + 1; i < n; ++i) { const double v = fabs(A[i * n + k]); if (v > best) { best = v; piv = i; } }
    if (piv != k) { det = -det; for (int j = k; j < n; ++j) { const double t = A[k * n + j]; A[k * n + j] = A[piv * n + j]; A[piv * n + j] = t; } }
    const double d = A[k * n + k];
    det *= d;
    if (d != 0.0) for (int i = k + 1; i < n; ++i) {
      const double f = A[i * n + k] / d;
      if (f != 0.0) for (int j = k + 1; j < n; ++j) A[i * n + j] -= f * A[k * n + j];
    }
  }
  return det;
}
STK_DEV int find_sorted_insert_index_v(double e, const VecD& v) {     // CC.hs:192-197
  int j = v.len + 1;
  for (int i = 1; i <= v.len; ++i) if (e < v.at(i)) { j = i; break; }
  return j;
}

// ---- sub-warp reductions ---------------------------------------------------
// digamma(x) = d/dx lgamma(x), x > 0 (the adjoint of lgamma in the gradient lowering: gamma / beta / inv_gamma densities whose
// shape is a parameter; stan_math uses boost::math::digamma): recurrence up to x >= 6, then the asymptotic series
// log x - 1/(2x) - sum B_2k / (2k x^2k) through x^-14 (truncation error < 2e-14 at x = 6)
STK_DEV double digamma(double x) {
  double r = 0.0;
  while (x < 6.0) { r -= 1.0 / x; x += 1.0; }
  const double f = 1.0 / (x * x);
  const double t = f * (-1.0 / 12.0 + f * (1.0 / 120.0 + f * (-1.0 / 252.0 + f * (1.0 / 240.0 + f * (-1.0 / 132.0 + f * (691.0 / 32760.0 + f * (-1.0 / 12.0)))))));
  return r + log(x) - 0.5 / x + t;
}

// sum over the G consecutive lanes of a chain's lane group (G = 1,2,4,...,32).
template <int G>
STK_DEV double group_sum(double v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// the same over an explicit lane mask (the G lanes of ONE chain's group): usable in divergent code, where other groups of the
// warp are elsewhere; every lane of the group ends with the same value (xor butterfly, fixed order)
template <int G>
STK_DEV double group_sum_masked(double v, unsigned mask) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o);
  return v;
}
template <int G>
STK_DEV int group_sum_i(int v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// record one double of a draw: draws[slot][r][chain]
STK_DEV void record_value(const StkArgs& a, long long slot, int r, long long chain, double v) {
  a.draws[(slot * a.n_rec + r) * a.n_chains + chain] = v;
}
// should transition t (0-based within the launch) be recorded, and where?
STK_DEV bool record_slot(const StkArgs& a, long long t, long long* slot) {
  if (!a.draws || a.draw_slots <= 0) return false;
  long long g = a.step0 + t + 1;           // transitions completed so far (global)
  if (g % a.draw_thin != 0) return false;
  // multiples of thin in (step0, g] since the launch started
  long long s = a.draw_base + (g / a.draw_thin - a.step0 / a.draw_thin) - 1;
  if (s >= a.draw_slots) return false;
  *slot = s;
  return true;
}

}  // namespace stk
#endif  // STK_DEVICE_CUH
