// stk_oracle_gp.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// Config C3, the form of gp.ipynb cells 6-15 (GP classifier with the squared-exponential kernel `kernelSE1`,
// hyper-parameters fixed, whitened latent `normalChol`, Program.hs:668-673), on the parameter space of the Stan
// program the notebook prints (cell 15):
//   K_ij = exp(-(x_i - x_j)^2 / 2) + (x_i == x_j ? 1e-6 : 0);  L = chol K (Eigen LLT in the reference, CC.hs:161-162);
//   g = L w;   lp(w) = sum_i normal_lpdf(w_i | 0, 1) + sum_i bernoulli_logit_lpmf(z_i | g_i)
//   grad = -w + L^T (z - sigmoid(g))
// PINNED by tests/golden/gp_stan.npz (vendored CmdStan 2.20 on that Stan text, log_prob_grad at 17 digits).
// HMC transition as orc_linreg_hmc_step (base_static_hmc.hpp:42-72, unit_e).
#include <atomic>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>
#include "stk_oracle.h"

extern "C" {

// lower Cholesky factor of the SE covariance of x[n], row-major [n][n] (upper part zero)
int orc_gp_chol(const double* x, int n, double* L) {
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      const double d = x[i] - x[j];
      L[(size_t)i * n + j] = j <= i ? std::exp(-(d * d) / 2.0) + (x[i] == x[j] ? 1.0e-6 : 0.0) : 0.0;
    }
  for (int j = 0; j < n; ++j) {
    double s = L[(size_t)j * n + j];
    for (int k = 0; k < j; ++k) s -= L[(size_t)j * n + k] * L[(size_t)j * n + k];
    if (!(s > 0.0)) return -1;
    const double d = std::sqrt(s);
    L[(size_t)j * n + j] = d;
    for (int i = j + 1; i < n; ++i) {
      double t = L[(size_t)i * n + j];
      for (int k = 0; k < j; ++k) t -= L[(size_t)i * n + k] * L[(size_t)j * n + k];
      L[(size_t)i * n + j] = t / d;
    }
  }
  return 0;
}

static double log1p_exp(double a) { return a > 0 ? a + std::log1p(std::exp(-a)) : std::log1p(std::exp(a)); }

double orc_gp_logp(const double* L, const int* z, int n, const double* w, double* grad) {
  std::vector<double> g(n), r(n);
  double lp = 0.0;
  for (int i = 0; i < n; ++i) lp += orc_normal_lpdf(w[i], 0, 1);
  for (int i = 0; i < n; ++i) {
    double s = 0.0;
    for (int k = 0; k <= i; ++k) s += L[(size_t)i * n + k] * w[k];
    g[i] = s;
    lp += z[i] ? -log1p_exp(-s) : -log1p_exp(s);                    // bernoulli_logit_lpmf (stan_math: log_inv_logit / log1m_inv_logit)
    r[i] = (double)z[i] - 1.0 / (1.0 + std::exp(-s));
  }
  if (grad)
    for (int k = 0; k < n; ++k) {
      double s = 0.0;
      for (int i = k; i < n; ++i) s += L[(size_t)i * n + k] * r[i];
      grad[k] = s - w[k];
    }
  return lp;
}

int orc_gp_hmc_step(const double* L, const int* z, int n, double* w, double eps, int Lf,
                    uint64_t seed, uint64_t chain, uint64_t step, double* lp_out) {
  orc_rng g; orc_rng_init(&g, seed, chain, step);
  std::vector<double> p(n + 1), gr(n), w0(w, w + n);
  for (int j = 0; j < n; j += 2) { double a, b; orc_rng_normal2(&g, &a, &b); p[j] = a; if (j + 1 < n) p[j + 1] = b; }
  double lp = orc_gp_logp(L, z, n, w, gr.data());
  const double lp0 = lp;
  double ke = 0; for (int j = 0; j < n; ++j) ke += p[j] * p[j];
  const double H0 = -lp + 0.5 * ke;
  for (int l = 0; l < Lf; ++l) {
    for (int j = 0; j < n; ++j) { p[j] += 0.5 * eps * gr[j]; w[j] += eps * p[j]; }
    lp = orc_gp_logp(L, z, n, w, gr.data());
    for (int j = 0; j < n; ++j) p[j] += 0.5 * eps * gr[j];
  }
  ke = 0; for (int j = 0; j < n; ++j) ke += p[j] * p[j];
  double H = -lp + 0.5 * ke;
  if (std::isnan(H)) H = std::numeric_limits<double>::infinity();
  const double accept_prob = std::exp(H0 - H);
  orc_rng gu; orc_rng_init(&gu, seed, chain, step); gu.slot = (uint32_t)((n + 1) / 2);
  const double u = orc_u01(&gu);
  int acc = 1;
  if (accept_prob < 1 && u > accept_prob) { std::memcpy(w, w0.data(), sizeof(double) * n); lp = lp0; acc = 0; }
  if (lp_out) *lp_out = lp;
  return acc;
}

// CPU-baseline driver: W is [n][C]
long long orc_gp_hmc_run(const double* L, const int* z, int n, double* W, long long C, long long T, double eps, int Lf,
                         uint64_t seed, uint64_t chain0, uint64_t step0, int nthreads) {
  std::atomic<long long> next{0}, accepted{0};
  auto work = [&]() {
    std::vector<double> wc(n);
    for (;;) {
      long long c = next.fetch_add(1);
      if (c >= C) break;
      for (int j = 0; j < n; ++j) wc[j] = W[(size_t)j * C + c];
      long long acc = 0;
      for (long long t = 0; t < T; ++t) acc += orc_gp_hmc_step(L, z, n, wc.data(), eps, Lf, seed, chain0 + c, step0 + t, nullptr);
      for (int j = 0; j < n; ++j) W[(size_t)j * C + c] = wc[j];
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
