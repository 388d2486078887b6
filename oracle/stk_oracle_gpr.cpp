// stk_oracle_gpr.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// Config C3 with PER-CHAIN hyper-parameters (SURVEY 8d "batched Cholesky per chain"): the GP marginal likelihood
//   lsv ~ normal 0 1;  lls2 ~ normal (log 100) 2                                  (gpclas priors, app/benchmark/Main.hs:84-91)
//   C = kernelSE' lsv lls2 x x = exp(lsv - distEucSq x x / (2 exp lls2)) + diag 0.01   (Main.hs:65-74)
//       distEucSq a b = asColumn (square a) + asRow (square b) - 2 *> outer a b        (the expanded form, kept as such)
//   y ~ multi_normal 0 C:  -0.5 * (|L^-1 y|^2 + n log 2 pi + LOG_DET(C))               (ccNode, CC.hs:221-225; LOG_DET :420-421)
// so one evaluation = covariance build + Cholesky (Eigen LLT in the reference, CC.hs:161-162) + one triangular solve.
// Transition: random-walk Metropolis on (lsv, lls2) in the shape of mh.ipynb cell 1 / orc_linreg_mh_step.
// PINNED by tests/golden/gpr_cc.json: the same statement templates compiled against the reference's vendored Eigen
// and Boost (oracle/ref_cc_mvn.cc -> oracle/_ref/cc_mvn), 17 digits.
#include <atomic>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>
#include "stk_oracle.h"

extern "C" {

double orc_gpr_logp(const double* x, const double* y, int n, const double* q) {
  const double lsv = q[0], lls2 = q[1];
  const double den = 2.0 * std::exp(lls2);
  std::vector<double> L((size_t)n * n), z(n);
  for (int i = 0; i < n; ++i)
    for (int j = 0; j <= i; ++j) {
      const double d2 = (x[i] * x[i] + x[j] * x[j]) - 2.0 * (x[i] * x[j]);
      L[(size_t)i * n + j] = std::exp(lsv - d2 / den) + (i == j ? 0.01 : 0.0);
    }
  double logdet = 0.0, quad = 0.0;
  for (int j = 0; j < n; ++j) {
    double s = L[(size_t)j * n + j];
    for (int k = 0; k < j; ++k) s -= L[(size_t)j * n + k] * L[(size_t)j * n + k];
    if (!(s > 0.0)) return std::nan("");
    const double d = std::sqrt(s);
    L[(size_t)j * n + j] = d;
    logdet += std::log(d);
    for (int i = j + 1; i < n; ++i) {
      double t = L[(size_t)i * n + j];
      for (int k = 0; k < j; ++k) t -= L[(size_t)i * n + k] * L[(size_t)j * n + k];
      L[(size_t)i * n + j] = t / d;
    }
    double t = y[j];
    for (int k = 0; k < j; ++k) t -= L[(size_t)j * n + k] * z[k];
    z[j] = t / d;
    quad += z[j] * z[j];
  }
  const double two_pi = 6.283185307179586476925286766559;
  double lp = -0.5 * (quad + std::log(two_pi) * n + 2.0 * logdet);
  lp += orc_normal_lpdf(lsv, 0.0, 1.0);
  lp += orc_normal_lpdf(lls2, std::log(100.0), 2.0);
  return lp;
}

// one random-walk Metropolis transition: slot 0 = the two proposal normals, slot 1 = the accept uniform
int orc_gpr_mh_step(const double* x, const double* y, int n, double* q, double rw, uint64_t seed, uint64_t chain,
                    uint64_t step, double* lp_io) {
  orc_rng g; orc_rng_init(&g, seed, chain, step);
  double a, b; orc_rng_normal2(&g, &a, &b);
  const double qn[2] = {q[0] + rw * a, q[1] + rw * b};
  const double lp = lp_io && !std::isnan(*lp_io) ? *lp_io : orc_gpr_logp(x, y, n, q);
  const double lpn = orc_gpr_logp(x, y, n, qn);
  const double u = orc_u01(&g);
  if (lpn == lpn && u < std::fmin(1.0, std::exp(lpn - lp))) { q[0] = qn[0]; q[1] = qn[1]; if (lp_io) *lp_io = lpn; return 1; }
  if (lp_io) *lp_io = lp;
  return 0;
}

// CPU-baseline / parity driver: Q is [2][C]; LP [C] (may be NULL) carries lp between calls (NaN = recompute)
long long orc_gpr_mh_run(const double* x, const double* y, int n, double* Q, double* LP, long long C, long long T, double rw,
                         uint64_t seed, uint64_t chain0, uint64_t step0, int nthreads) {
  std::atomic<long long> next{0}, accepted{0};
  auto work = [&]() {
    for (;;) {
      long long c = next.fetch_add(1);
      if (c >= C) break;
      double q[2] = {Q[c], Q[C + c]};
      double lp = LP ? LP[c] : std::nan("");
      long long acc = 0;
      for (long long t = 0; t < T; ++t) acc += orc_gpr_mh_step(x, y, n, q, rw, seed, chain0 + c, step0 + t, &lp);
      Q[c] = q[0]; Q[C + c] = q[1];
      if (LP) LP[c] = lp;
      accepted.fetch_add(acc);
    }
  };
  if (nthreads <= 1) { work(); return accepted.load(); }
  std::vector<std::thread> th;
  for (int i = 0; i < nthreads; ++i) th.emplace_back(work);
  for (auto& t : th) t.join();
  return accepted.load();
}

}  // extern "C"
