// stk_oracle_birats.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// Config C4: the hierarchical BiRats growth model (birats.ipynb cells 1-2, app/benchmark/Main.hs:106-125) on the
// unconstrained scale of the Stan program that Language.Stan generates for it (birats.ipynb cell 8):
//   parameters  vector[2] betaMu; vector[2] tau; corr_matrix[2] Omega; matrix[N,2] beta; real<lower=0> sigma
//   q = (betaMu[2], tau[2], y_omega, beta[.,1] (N), beta[.,2] (N), log sigma)            P = 2N + 6  (66 for N = 30)
//   lp = sum normal_lpdf(betaMu_k | 0, 100)
//      + sum cauchy_lpdf(tau_k | 0, 2.5) + log 2            [T[0,]: -inf when tau_k < 0]
//      + lkj_corr_lpdf(Omega | 2) + log(1 - z^2)            [z = tanh(y_omega): corr_constrain.hpp:45-47; read_corr_L adds 0 for K = 2]
//      + sum_i multi_normal_lpdf(beta_i | betaMu, diag(tau) Omega diag(tau))
//      + cauchy_lpdf(sigma | 0, 2.5) + log 2 + log sigma    [lb_constrain Jacobian]
//      + sum_ij normal_lpdf(y_ij | beta_i1 + beta_i2 x_j, sigma)
// with every additive constant kept (log_prob_grad<propto=false, jacobian=true>).
// PINNED by tests/golden/birats_stan.npz: lp and gradient from the vendored CmdStan 2.20 compiled on the Stan text
// of the notebook (oracle/stan/make_golden.py birats), and the posterior moments of its static-HMC sampler.
// HMC transition: base_static_hmc.hpp:42-72 + expl_leapfrog.hpp:17-34, unit_e metric (as orc_linreg_hmc_step).
#include <atomic>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>
#include "stk_oracle.h"

static const double kLogPi = 1.1447298858494001741434273513531;
static const double kLog2Pi = 1.8378770664093454835606594728112;

extern "C" double orc_birats_logp(const double* x, const double* y, int N, int Tn, const double* q, double* grad) {
  const int P = 2 * N + 6;
  const double mu1 = q[0], mu2 = q[1], t1 = q[2], t2 = q[3], yo = q[4], ls = q[P - 1];
  const double* b1 = q + 5; const double* b2 = q + 5 + N;
  if (grad) for (int j = 0; j < P; ++j) grad[j] = 0.0;
  if (t1 < 0.0 || t2 < 0.0) return -std::numeric_limits<double>::infinity();
  const double sigma = std::exp(ls), is2 = std::exp(-2.0 * ls);
  const double z = std::tanh(yo), D = 1.0 - z * z, lD = std::log1p(-z * z);
  double lp = orc_normal_lpdf(mu1, 0, 100) + orc_normal_lpdf(mu2, 0, 100);
  lp += (orc_cauchy_lpdf(t1, 0, 2.5) + std::log(2.0)) + (orc_cauchy_lpdf(t2, 0, 2.5) + std::log(2.0));
  // do_lkj_constant(eta = 2, K = 2) = lgamma(2.5) - (0.5 log pi + lgamma(2))   (lkj_corr_lpdf.hpp)
  lp += (std::lgamma(2.5) - (0.5 * kLogPi + std::lgamma(2.0))) + (2.0 - 1.0) * lD;
  lp += lD;                                                            // Jacobian of tanh
  lp += (orc_cauchy_lpdf(sigma, 0, 2.5) + std::log(2.0)) + ls;
  double mvn = 0.0, lik = 0.0, ss = 0.0;
  double g_mu1 = 0, g_mu2 = 0, g_t1 = 0, g_t2 = 0, g_z = 0;
  for (int i = 0; i < N; ++i) {
    const double a = (b1[i] - mu1) / t1, c = (b2[i] - mu2) / t2;
    const double Q = (a * a - 2.0 * z * a * c + c * c) / D;
    mvn += -kLog2Pi - std::log(t1) - std::log(t2) - 0.5 * lD - 0.5 * Q;
    const double da = (a - z * c) / D, dc = (c - z * a) / D;             // dQ/da / 2, dQ/dc / 2
    double sr = 0.0, srx = 0.0;
    for (int j = 0; j < Tn; ++j) {
      const double m = b1[i] + b2[i] * x[j], r = y[i * Tn + j] - m;
      lik += orc_normal_lpdf(y[i * Tn + j], m, sigma);
      ss += r * r; sr += r; srx += r * x[j];
    }
    if (grad) {
      grad[5 + i] = -da / t1 + sr * is2;
      grad[5 + N + i] = -dc / t2 + srx * is2;
      g_mu1 += da / t1; g_mu2 += dc / t2;
      g_t1 += -1.0 / t1 + a * da / t1; g_t2 += -1.0 / t2 + c * dc / t2;
      g_z += z / D - 0.5 * ((-2.0 * a * c + 2.0 * z * Q) / D);
    }
  }
  lp += mvn + lik;
  if (grad) {
    grad[0] = g_mu1 - mu1 / 1e4; grad[1] = g_mu2 - mu2 / 1e4;
    grad[2] = g_t1 - 2.0 * t1 / (6.25 + t1 * t1); grad[3] = g_t2 - 2.0 * t2 / (6.25 + t2 * t2);
    grad[4] = (g_z - 4.0 * z / D) * D;                                  // (lkj + Jacobian: 2 log(1 - z^2)) ; dz/dy = 1 - z^2
    const double zc = sigma / 2.5;
    grad[P - 1] = -(double)(N * Tn) + ss * is2 - 2.0 * zc * zc / (1.0 + zc * zc) + 1.0;
  }
  return lp;
}

extern "C" int orc_birats_hmc_step(const double* x, const double* y, int N, int Tn, double* q, double eps, int L,
                                   uint64_t seed, uint64_t chain, uint64_t step, double* lp_out) {
  const int P = 2 * N + 6;
  orc_rng g; orc_rng_init(&g, seed, chain, step);
  std::vector<double> p(P + 1), gr(P), q0(q, q + P);
  for (int j = 0; j < P; j += 2) { double a, b; orc_rng_normal2(&g, &a, &b); p[j] = a; if (j + 1 < P) p[j + 1] = b; }
  double lp = orc_birats_logp(x, y, N, Tn, q, gr.data());
  const double lp0 = lp;
  double ke = 0; for (int j = 0; j < P; ++j) ke += p[j] * p[j];
  const double H0 = -lp + 0.5 * ke;
  for (int l = 0; l < L; ++l) {
    for (int j = 0; j < P; ++j) { p[j] += 0.5 * eps * gr[j]; q[j] += eps * p[j]; }
    lp = orc_birats_logp(x, y, N, Tn, q, gr.data());
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

extern "C" long long orc_birats_hmc_run(const double* x, const double* y, int N, int Tn, double* q, long long C,
                                        long long T, double eps, int L, uint64_t seed, uint64_t chain0,
                                        uint64_t step0, int nthreads) {
  const int P = 2 * N + 6;
  std::atomic<long long> next{0}, accepted{0};
  auto work = [&]() {
    std::vector<double> qc(P);
    for (;;) {
      long long c = next.fetch_add(1);
      if (c >= C) break;
      for (int j = 0; j < P; ++j) qc[j] = q[(size_t)j * C + c];
      long long acc = 0;
      for (long long t = 0; t < T; ++t) acc += orc_birats_hmc_step(x, y, N, Tn, qc.data(), eps, L, seed, chain0 + c, step0 + t, nullptr);
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
