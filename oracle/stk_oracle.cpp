// stk_oracle.cpp — CPU ORACLE, test infrastructure only (see stk_oracle.h).
// Sequential, scalar, obviously-correct restatements; sums run in the
// reference's order (left fold over 1..n, CC.hs:236-262 / Program.hs:858-865).
#include "stk_oracle.h"

#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

static const double kHalfLog2Pi = 0.91893853320467274178032973640562;
static const double kLogPi = 1.1447298858494001741434273513531;
static const double kPi = 3.14159265358979323846264338327950288;

// ---------------------------------------------------------------------------
// Philox4x32-10, Salmon et al. SC'11 (Random123 reference constants)
// ---------------------------------------------------------------------------
extern "C" void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c0 = n0; c1 = (uint32_t)p1; c2 = n2; c3 = (uint32_t)p0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static double u01_from_bits(uint32_t hi, uint32_t lo) {
  uint64_t m = ((uint64_t)(hi >> 5) << 26) | (uint64_t)(lo >> 6);
  return ((double)m + 0.5) * 1.1102230246251565404236316680908e-16;
}

extern "C" void orc_rng_init(orc_rng* g, uint64_t seed, uint64_t chain, uint64_t step) {
  g->k0 = (uint32_t)seed; g->k1 = (uint32_t)(seed >> 32);
  g->c0 = (uint32_t)chain; g->c1 = (uint32_t)step; g->c2 = (uint32_t)(step >> 32);
  g->slot = 0;
}
static void rng_next(orc_rng* g, uint32_t o[4]) {
  uint32_t ctr[4] = {g->c0, g->c1, g->c2, g->slot++}, key[2] = {g->k0, g->k1};
  orc_philox4x32_10(ctr, key, o);
}
static void rng_u01x2(orc_rng* g, double* a, double* b) {
  uint32_t o[4]; rng_next(g, o);
  *a = u01_from_bits(o[0], o[1]); *b = u01_from_bits(o[2], o[3]);
}
static void rng_normal2(orc_rng* g, double* a, double* b) {
  double u1, u2; rng_u01x2(g, &u1, &u2);
  double r = std::sqrt(-2.0 * std::log(u1));
  *a = r * std::cos(2.0 * kPi * u2); *b = r * std::sin(2.0 * kPi * u2);
}
extern "C" double orc_u01(orc_rng* g) { double a, b; rng_u01x2(g, &a, &b); return a; }
extern "C" void orc_rng_normal2(orc_rng* g, double* a, double* b) { rng_normal2(g, a, b); }
extern "C" double orc_normal(orc_rng* g) { double a, b; rng_normal2(g, &a, &b); return a; }
extern "C" void orc_u01x2(uint64_t seed, uint64_t chain, uint64_t step, uint32_t slot, double* a, double* b) {
  orc_rng g; orc_rng_init(&g, seed, chain, step); g.slot = slot; rng_u01x2(&g, a, b);
}
extern "C" void orc_normal2(uint64_t seed, uint64_t chain, uint64_t step, uint32_t slot, double* a, double* b) {
  orc_rng g; orc_rng_init(&g, seed, chain, step); g.slot = slot; rng_normal2(&g, a, b);
}
extern "C" int orc_draw_discrete_uniform(orc_rng* g, int a, int b) {
  int n = b - a + 1;
  int k = (int)std::floor(orc_u01(g) * (double)n);
  if (k >= n) k = n - 1;
  return a + k;
}
extern "C" int orc_draw_bernoulli(orc_rng* g, double p) { return orc_u01(g) < p ? 1 : 0; }
extern "C" int orc_draw_pmf(orc_rng* g, const double* ps, int n) {
  double tot = 0.0;
  for (int i = 0; i < n; ++i) tot += ps[i];
  double u = orc_u01(g) * tot, acc = 0.0;
  for (int i = 0; i < n; ++i) { acc += ps[i]; if (u < acc) return i + 1; }
  return n;
}
static int poisson_small(orc_rng* g, double l) {
  double u = orc_u01(g);
  double p = std::exp(-l), c = p;
  int k = 0;
  while (u >= c && k < 1000) { ++k; p *= l / (double)k; c += p; }
  return k;
}
extern "C" int orc_draw_poisson(orc_rng* g, double l) {
  if (l < 30.0) return poisson_small(g, l);
  int m = (int)(l / 25.0) + 1, tot = 0;
  for (int j = 0; j < m; ++j) tot += poisson_small(g, l / (double)m);
  return tot;
}
extern "C" double orc_draw_gamma(orc_rng* g, double a, double b) {
  double boost = 1.0;
  if (a < 1.0) { boost = std::pow(orc_u01(g), 1.0 / a); a += 1.0; }
  double d = a - 1.0 / 3.0, c = 1.0 / std::sqrt(9.0 * d);
  for (int it = 0; it < 64; ++it) {
    double x = orc_normal(g);
    double v = 1.0 + c * x;
    if (v <= 0.0) continue;
    v = v * v * v;
    double u = orc_u01(g);
    if (std::log(u) < 0.5 * x * x + d - d * v + d * std::log(v)) return boost * d * v / b;
  }
  return boost * d / b;
}
extern "C" double orc_draw_cauchy(orc_rng* g, double m, double s) {
  double x, y; rng_normal2(g, &x, &y); return m + s * (x / y);
}

extern "C" double orc_draw_inv_gamma(orc_rng* g, double a, double b) { return 1.0 / orc_draw_gamma(g, a, b); }  // Abstract.hs:72-75
extern "C" int orc_draw_geometric(orc_rng* g, double p) {  // Abstract.hs:80-85 (the law; one uniform by inversion)
  const double u = orc_u01(g);
  if (p >= 1.0) return 0;
  const double k = std::floor(std::log(u) / std::log1p(-p));
  return k < 2147483647.0 ? (int)k : 2147483647;
}
extern "C" double orc_draw_logistic(orc_rng* g, double m, double s) { const double u = orc_u01(g); return m + s * std::log(u / (1.0 - u)); }
extern "C" int orc_draw_neg_binomial(orc_rng* g, double a, double b) { const double l = orc_draw_gamma(g, a, b); return orc_draw_poisson(g, l); }
extern "C" double orc_draw_beta(orc_rng* g, double a, double b) {
  const double X = orc_draw_gamma(g, a, 1.0), Y = orc_draw_gamma(g, b, 1.0);
  return X / (X + Y);
}
extern "C" double orc_draw_lkj_corr2(orc_rng* g, double v) { return 2.0 * orc_draw_beta(g, v, v) - 1.0; }

// ---------------------------------------------------------------------------
// log densities
// ---------------------------------------------------------------------------
extern "C" double orc_log_factorial(int k) { return std::lgamma(1.0 + (double)k); }  // CC.hs:167-168
extern "C" double orc_normal_lpdf(double x, double m, double s) {  // Abstract.hs:121
  return -std::log(2 * kPi) / 2 - std::log(s) - std::pow((x - m) / s, 2) / 2;
}
extern "C" double orc_gamma_lpdf(double x, double a, double b) {  // Abstract.hs:68
  return a * std::log(b) + (a - 1) * std::log(x) - b * x - std::lgamma(a);
}
extern "C" double orc_inv_gamma_lpdf(double x, double a, double b) {  // Abstract.hs:77
  return a * std::log(b) + (-a - 1) * std::log(x) - b / x - std::lgamma(a);
}
extern "C" double orc_geometric_lpdf(int k, double p) { return k * std::log(1 - p) + std::log(p); }  // :90
extern "C" double orc_neg_binomial_lpdf(int k, double a, double b) {  // Abstract.hs:113-115
  return std::lgamma(a + k) - std::lgamma(a) - orc_log_factorial(k) + k * std::log(1 / (b + 1)) + a * std::log(b / (b + 1));
}
extern "C" double orc_poisson_lpdf(int k, double l) { return k * std::log(l) - orc_log_factorial(k) - l; }  // :152
extern "C" double orc_uniform_lpdf(double x, double a, double b) {  // :161-162
  if (x < a || b < x) return -std::numeric_limits<double>::infinity();
  return std::log(1 / (b - a));
}
extern "C" double orc_uniform_cdf(double x, double a, double b) {  // :163-165
  if (x < a) return 0;
  if (x > b) return 1;
  return (x - a) / (b - a);
}
extern "C" double orc_discrete_uniform_lpdf(int x, int a, int b) {  // :166-167
  if (x < a || b < x) return -std::numeric_limits<double>::infinity();
  return std::log(1.0 / (double)(b - a + 1));
}
extern "C" double orc_bernoulli_logit_lpdf(int b, double l) {  // Const.hs:92-93
  double p = 1 / (1 + std::exp(-l));
  return std::log(b ? p : 1 - p);
}
extern "C" double orc_cauchy_lpdf(double x, double m, double s) {  // stan_math cauchy_lpdf
  double z = (x - m) / s;
  return -kLogPi - std::log(s) - std::log1p(z * z);
}

// ---------------------------------------------------------------------------
// C1 linreg.  lp = sum of the PNode densities of linreg.ipynb cell 1 (d-dim w)
// on the unconstrained scale q = (b, w, log sigma):
//   normal_lpdf(b;0,10) + sum_j normal_lpdf(w_j;0,10)
//   + cauchy_lpdf(sigma;0,5) - log(1 - cauchy_cdf(0;0,5))     [truncated 0 inf]
//   + sum_i normal_lpdf(y_i; b + x_i.w, sigma)                 [normals, left fold]
//   + log sigma                                                 [lb_constrain Jacobian]
// ---------------------------------------------------------------------------
extern "C" double orc_linreg_logp(const double* X, const double* y, int n, int d, const double* q, double* grad) {
  const double b = q[0];
  const double* w = q + 1;
  const double ls = q[d + 1], sigma = std::exp(ls);
  double lp = orc_normal_lpdf(b, 0, 10);
  for (int j = 0; j < d; ++j) lp += orc_normal_lpdf(w[j], 0, 10);
  lp += orc_cauchy_lpdf(sigma, 0, 5) + std::log(2.0);
  std::vector<double> gw(d, 0.0);
  double gb = 0.0, ss = 0.0, lik = 0.0;
  for (int i = 0; i < n; ++i) {
    double mu = b;
    for (int j = 0; j < d; ++j) mu += X[(size_t)i * d + j] * w[j];
    double r = y[i] - mu;
    lik += orc_normal_lpdf(y[i], mu, sigma);
    ss += r * r; gb += r;
    for (int j = 0; j < d; ++j) gw[j] += X[(size_t)i * d + j] * r;
  }
  lp += lik + ls;
  if (grad) {
    double is2 = 1.0 / (sigma * sigma);
    grad[0] = gb * is2 - b / 100.0;
    for (int j = 0; j < d; ++j) grad[1 + j] = gw[j] * is2 - w[j] / 100.0;
    double z2 = (sigma / 5.0) * (sigma / 5.0);
    grad[d + 1] = -(double)n + ss * is2 - 2.0 * z2 / (1.0 + z2) + 1.0;
  }
  return lp;
}

extern "C" int orc_linreg_hmc_step(const double* X, const double* y, int n, int d, double* q, double eps, int L,
                                   uint64_t seed, uint64_t chain, uint64_t step, double* lp_out) {
  const int P = d + 2;
  orc_rng g; orc_rng_init(&g, seed, chain, step);
  std::vector<double> p(P + 1), gr(P), q0(q, q + P);
  for (int j = 0; j < P; j += 2) { double a, b; rng_normal2(&g, &a, &b); p[j] = a; if (j + 1 < P) p[j + 1] = b; }
  double lp = orc_linreg_logp(X, y, n, d, q, gr.data());
  const double lp0 = lp;
  double ke = 0; for (int j = 0; j < P; ++j) ke += p[j] * p[j];
  const double H0 = -lp + 0.5 * ke;
  for (int l = 0; l < L; ++l) {
    for (int j = 0; j < P; ++j) { p[j] += 0.5 * eps * gr[j]; q[j] += eps * p[j]; }
    lp = orc_linreg_logp(X, y, n, d, q, gr.data());
    for (int j = 0; j < P; ++j) p[j] += 0.5 * eps * gr[j];
  }
  ke = 0; for (int j = 0; j < P; ++j) ke += p[j] * p[j];
  double H = -lp + 0.5 * ke;
  if (std::isnan(H)) H = std::numeric_limits<double>::infinity();
  const double accept_prob = std::exp(H0 - H);
  const double u = orc_u01(&g);
  int acc = 1;
  if (accept_prob < 1 && u > accept_prob) { std::memcpy(q, q0.data(), sizeof(double) * P); lp = lp0; acc = 0; }
  if (lp_out) *lp_out = lp;
  return acc;
}

extern "C" int orc_linreg_mh_step(const double* X, const double* y, int n, int d, double* q, double rw,
                                  uint64_t seed, uint64_t chain, uint64_t step, double* lp_io) {
  const int P = d + 2;
  orc_rng g; orc_rng_init(&g, seed, chain, step);
  std::vector<double> qn(P + 1);
  for (int j = 0; j < P; j += 2) { double a, b; rng_normal2(&g, &a, &b); qn[j] = q[j] + rw * a; if (j + 1 < P) qn[j + 1] = q[j + 1] + rw * b; }
  const double lp = lp_io && !std::isnan(*lp_io) ? *lp_io : orc_linreg_logp(X, y, n, d, q, nullptr);
  const double lpn = orc_linreg_logp(X, y, n, d, qn.data(), nullptr);
  const double u = orc_u01(&g);
  if (u < std::fmin(1.0, std::exp(lpn - lp))) { std::memcpy(q, qn.data(), sizeof(double) * P); if (lp_io) *lp_io = lpn; return 1; }
  if (lp_io) *lp_io = lp;
  return 0;
}

extern "C" void orc_linreg_simulate(const double* X, int n, int d, uint64_t seed, uint64_t chain, double* q, double* ysim) {
  orc_rng g; orc_rng_init(&g, seed, chain, ~0ULL);
  for (int j = 0; j < d + 1; j += 2) { double a, b; rng_normal2(&g, &a, &b); q[j] = 10 * a; if (j + 1 < d + 1) q[j + 1] = 10 * b; }
  double sigma = 1.0;
  for (int it = 0; it < 64; ++it) { sigma = orc_draw_cauchy(&g, 0, 5); if (sigma >= 0) break; }
  sigma = std::fabs(sigma);
  q[d + 1] = std::log(sigma);
  for (int i = 0; i < n; ++i) {
    double mu = q[0];
    for (int j = 0; j < d; ++j) mu += X[(size_t)i * d + j] * q[1 + j];
    orc_rng g2 = g; g2.slot = 64u + (uint32_t)i;
    ysim[i] = mu + sigma * orc_normal(&g2);
  }
}

// ---------------------------------------------------------------------------
// CPU-baseline driver (bench.py cpu_baseline / --impl reference): C chains x T
// static-HMC transitions, one chain at a time per host thread (the reference's
// own execution shape: one sequential chain per process/core).  q is [P][C].
// ---------------------------------------------------------------------------
#include <atomic>
#include <thread>
extern "C" long long orc_linreg_hmc_run(const double* X, const double* y, int n, int d, double* q, long long C,
                                        long long T, double eps, int L, uint64_t seed, uint64_t chain0,
                                        uint64_t step0, int nthreads) {
  const int P = d + 2;
  std::atomic<long long> next{0}, accepted{0};
  auto work = [&]() {
    std::vector<double> qc(P);
    for (;;) {
      long long c = next.fetch_add(1);
      if (c >= C) break;
      for (int j = 0; j < P; ++j) qc[j] = q[(size_t)j * C + c];
      long long acc = 0;
      for (long long t = 0; t < T; ++t) acc += orc_linreg_hmc_step(X, y, n, d, qc.data(), eps, L, seed, chain0 + c, step0 + t, nullptr);
      for (int j = 0; j < P; ++j) q[(size_t)j * C + c] = qc[j];
      accepted.fetch_add(acc);
    }
  };
  if (nthreads <= 1) { work(); return accepted.load(); }
  std::vector<std::thread> th;
  for (int i = 0; i < nthreads; ++i) th.emplace_back(work);
  for (auto& t : th) t.join();
  return accepted.load();
}
