// stk_oracle_dp.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// Config C5: Dirichlet-process mixture  dpmm n  (dp.ipynb cells 1-3, app/dp/Main.hs:8-31):
//   sticks_k ~ beta 1 alpha;  probs_k = sticks_k * prod_{l<k} (1 - sticks_l);  atoms_k ~ uniform 0 100;
//   params_j = atoms[stick_j], stick_j ~ pmf probs;  values_j ~ normal params_j 1,        alpha = 5
// The reference can only FORWARD-SIMULATE this model (Church backend, lazily memoised infinite sticks); it has no
// inference for it.  PARITY UNPINNED: this file is our restatement of (a) the forward model truncated at K sticks and
// (b) the standard blocked Gibbs sampler for the truncated stick-breaking representation (Ishwaran & James 2001),
// driven by the same Philox stream as the CUDA kernels (models/dp.cu):
//   counter = (chain, sweep, slot):  slot i/2, component i%2 -> u_i of point i;
//   slot 2^30 + 64 k (+32) -> the two gamma draws of stick k;  slot 2^30 + 2^16 + 64 k -> atom k.
// Sufficient statistics are exact: n_k integer, sum_k y_i accumulated as int64 of round(y_i * 2^36) — the result does
// not depend on the order of accumulation, so the parallel kernel and this loop agree bit for bit.
#include <cmath>
#include <cstring>
#include <thread>
#include <atomic>
#include <vector>
#include "stk_oracle.h"

static const uint32_t SB = 1u << 30, AB = (1u << 30) + (1u << 16);
static const double FIX = 68719476736.0;   // 2^36

// assignment step for one point: returns the 1-based cluster
static int dp_assign(double y, const double* lpi, const double* mu, int K, double u) {
  double lw[64], m = -INFINITY;
  for (int k = 0; k < K; ++k) { const double d = y - mu[k]; lw[k] = std::fma(-0.5 * d, d, lpi[k]); if (lw[k] > m) m = lw[k]; }
  double S = 0.0, w[64];
  for (int k = 0; k < K; ++k) { const double t = lw[k] - m; w[k] = (t < -50.0) ? 0.0 : std::exp(t); S += w[k]; }
  const double target = u * S;
  double cum = 0.0; int z = K;
  for (int k = 0; k < K; ++k) { cum += w[k]; if (target < cum) { z = k + 1; break; } }
  return z;
}

static void dp_logpi(const double* v, int K, double* lpi) {
  double rem = 1.0;
  for (int k = 0; k < K; ++k) { const double vk = (k == K - 1) ? 1.0 : v[k]; lpi[k] = std::log(vk * rem); rem *= (1.0 - vk); }
}

extern "C" {

// one blocked-Gibbs sweep of one chain; v[K], mu[K] updated in place, z[N] written (1-based)
void orc_dp_sweep(const double* y, int N, int K, double alpha, double lo, double hi, double* v, double* mu, int* z,
                  uint64_t seed, uint64_t chain, uint64_t sweep) {
  std::vector<double> lpi(K);
  dp_logpi(v, K, lpi.data());
  std::vector<long long> cnt(K, 0), sum(K, 0);
  for (int i = 0; i < N; i += 2) {
    double u0, u1;
    orc_u01x2(seed, chain, sweep, (uint32_t)(i >> 1), &u0, &u1);
    for (int e = 0; e < 2 && i + e < N; ++e) {
      const int zi = dp_assign(y[i + e], lpi.data(), mu, K, e ? u1 : u0);
      z[i + e] = zi; cnt[zi - 1] += 1; sum[zi - 1] += std::llrint(y[i + e] * FIX);
    }
  }
  long long after = 0;
  std::vector<long long> suffix(K, 0);
  for (int k = K - 1; k >= 0; --k) { suffix[k] = after; after += cnt[k]; }
  for (int k = 0; k < K; ++k) {
    if (k < K - 1) {
      orc_rng g; orc_rng_init(&g, seed, chain, sweep); g.slot = SB + 64u * (uint32_t)k;
      const double X = orc_draw_gamma(&g, 1.0 + (double)cnt[k], 1.0);
      g.slot = SB + 64u * (uint32_t)k + 32u;
      const double Y = orc_draw_gamma(&g, alpha + (double)suffix[k], 1.0);
      v[k] = X / (X + Y);
    } else v[k] = 1.0;
    orc_rng g; orc_rng_init(&g, seed, chain, sweep); g.slot = AB + 64u * (uint32_t)k;
    if (cnt[k] > 0) {
      const double mean = ((double)sum[k] / FIX) / (double)cnt[k], sd = 1.0 / std::sqrt((double)cnt[k]);
      double m = mean;
      for (int it = 0; it < 32; ++it) { m = mean + sd * orc_normal(&g); if (m > lo && m < hi) break; }
      mu[k] = std::fmin(std::fmax(m, lo), hi);
    } else mu[k] = lo + (hi - lo) * orc_u01(&g);
  }
}

// forward simulation of dpmm n truncated at K sticks (sweep counter 2^64-1): v, mu from the prior, then per point
// stick_j ~ pmf probs (slot j/2) and values_j = atoms[stick_j] + normal (slot 2^29 + j)
void orc_dp_simulate(int N, int K, double alpha, double lo, double hi, uint64_t seed, uint64_t chain,
                     double* v, double* mu, int* z, double* ysim) {
  const uint64_t T = ~0ULL;
  for (int k = 0; k < K; ++k) {
    orc_rng g; orc_rng_init(&g, seed, chain, T); g.slot = SB + 64u * (uint32_t)k;
    const double X = orc_draw_gamma(&g, 1.0, 1.0);
    g.slot = SB + 64u * (uint32_t)k + 32u;
    const double Y = orc_draw_gamma(&g, alpha, 1.0);
    v[k] = (k == K - 1) ? 1.0 : X / (X + Y);
    g.slot = AB + 64u * (uint32_t)k;
    mu[k] = lo + (hi - lo) * orc_u01(&g);
  }
  std::vector<double> pr(K);
  double rem = 1.0;
  for (int k = 0; k < K; ++k) { pr[k] = v[k] * rem; rem *= (1.0 - v[k]); }
  for (int i = 0; i < N; ++i) {
    double u0, u1; orc_u01x2(seed, chain, T, (uint32_t)(i >> 1), &u0, &u1);
    const double u = (i & 1) ? u1 : u0;
    double cum = 0.0; int zi = K;
    for (int k = 0; k < K; ++k) { cum += pr[k]; if (u < cum) { zi = k + 1; break; } }   // Church `pmf`: first j with u <= cumulative
    z[i] = zi;
    double a, b; orc_normal2(seed, chain, T, (1u << 29) + (uint32_t)i, &a, &b);
    ysim[i] = mu[zi - 1] + a;
  }
}

// CPU-baseline driver: C chains x T sweeps; v, mu are [K][C]; z is not kept (the baseline measures the sweep)
long long orc_dp_run(const double* y, int N, int K, double alpha, double lo, double hi, double* v, double* mu,
                     long long C, long long T, uint64_t seed, uint64_t chain0, uint64_t step0, int nthreads) {
  std::atomic<long long> next{0};
  auto work = [&]() {
    std::vector<double> vc(K), mc(K); std::vector<int> z(N);
    for (;;) {
      long long c = next.fetch_add(1);
      if (c >= C) break;
      for (int k = 0; k < K; ++k) { vc[k] = v[(size_t)k * C + c]; mc[k] = mu[(size_t)k * C + c]; }
      for (long long t = 0; t < T; ++t) orc_dp_sweep(y, N, K, alpha, lo, hi, vc.data(), mc.data(), z.data(), seed, chain0 + c, step0 + t);
      for (int k = 0; k < K; ++k) { v[(size_t)k * C + c] = vc[k]; mu[(size_t)k * C + c] = mc[k]; }
    }
  };
  if (nthreads <= 1) { work(); return C * T; }
  std::vector<std::thread> th;
  for (int i = 0; i < nthreads; ++i) th.emplace_back(work);
  for (auto& t : th) t.join();
  return C * T;
}

}  // extern "C"
