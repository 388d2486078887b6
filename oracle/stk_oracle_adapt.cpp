// stk_oracle_adapt.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// Restatement of Stan's adaptive static HMC with a diagonal metric, the sampler behind
// `hmcStan`'s adaptation (Stan.hs:428-440 passes `adapt engaged=1`; static engine: app/benchmark/Main.hs:250-258):
//   * transition           cmdstan/stan/src/stan/mcmc/hmc/static/base_static_hmc.hpp:42-72
//   * diag_e metric        .../hmc/hamiltonians/diag_e_metric.hpp:20-60 (T = 0.5 p.(Minv p), p_i = z_i / sqrt(Minv_i))
//   * leapfrog             .../hmc/integrators/base_leapfrog.hpp:18-28, expl_leapfrog.hpp:17-34
//   * init_stepsize        .../hmc/base_hmc.hpp:73-126 (doubling / halving until the one-step acceptance crosses 0.8)
//   * dual averaging       .../stepsize_adaptation.hpp:67-87 (learn_stepsize, complete_adaptation)
//   * windows              .../windowed_adaptation.hpp:27-118, var_adaptation.hpp:20-40 (Welford + shrinkage)
//   * driver order         .../hmc/static/adapt_diag_e_static_hmc.hpp:29-51, services/util/run_adaptive_sampler.hpp:50-90,
//                          services/sample/hmc_static_diag_e_adapt.hpp:62-78 (mu = log(10 * stepsize) BEFORE init_stepsize)
// Random numbers come from the repo's Philox stream (seed, chain, transition) in the order the CUDA kernels use, so the
// trajectories of the kernels can be compared step by step; parity with CmdStan itself is statistical (adapted step
// size and metric, tests/golden/linreg_adapt_stan.json).
//
// Adaptation state per chain, `ad[ORC_AD_LEN]`:
//   0 eps (nominal)  1 mu  2 s_bar  3 x_bar  4 counter  5 window counter  6 next window end  7 window size
//   8 Welford n      9 phase (0 = not initialised, 1 = adapting, 2 = adaptation complete)
// Configuration, `cfg[ORC_CFG_LEN]`:
//   0 stepsize  1 int_time  2 num_warmup  3 delta  4 gamma  5 kappa  6 t0  7 init_buffer  8 term_buffer  9 window
#include "stk_oracle.h"
#include <cmath>
#include <cstring>
#include <functional>
#include <limits>
#include <vector>

namespace {
typedef std::function<double(const double* q, double* grad)> LogpGrad;

struct Windows { double num_warmup, init_buffer, term_buffer, base; bool windows_on; };

// windowed_adaptation::set_window_params.  Two reference quirks are kept: num_warmup < 20 leaves every member 0 (no
// estimation); and when the three stages do not fit, the 15/75/10 % split is stored but restart() is NOT called, so
// the first window end stays at its constructor value 0 + 0 - 1 (unsigned wrap): the metric is never updated.
Windows windows_of(const double* cfg) {
  Windows w;
  const double nw = std::floor(cfg[2]), ib = std::floor(cfg[7]), tb = std::floor(cfg[8]), bw = std::floor(cfg[9]);
  if (nw < 20) { w = {0, 0, 0, 0, false}; return w; }
  if (ib + bw + tb > nw) {
    const double ib2 = std::floor(0.15 * nw), tb2 = std::floor(0.10 * nw);
    w = {nw, ib2, tb2, nw - (ib2 + tb2), false};
    return w;
  }
  w = {nw, ib, tb, bw, true};
  return w;
}

struct Hmc {
  int P; const LogpGrad& f; orc_rng* g;
  std::vector<double> p, gr;
  Hmc(int P_, const LogpGrad& f_, orc_rng* g_) : P(P_), f(f_), g(g_), p(P_ + 1), gr(P_) {}
  void sample_p(const double* minv) {
    for (int j = 0; j < P; j += 2) {
      double a, b; orc_rng_normal2(g, &a, &b);
      p[j] = a / std::sqrt(minv[j]); if (j + 1 < P) p[j + 1] = b / std::sqrt(minv[j + 1]);
    }
  }
  double kinetic(const double* minv) const { double k = 0; for (int j = 0; j < P; ++j) k += minv[j] * p[j] * p[j]; return 0.5 * k; }
  // one leapfrog step from (q, p) with the gradient of lp at q in gr; returns lp at the new q
  double evolve(double* q, const double* minv, double eps) {
    for (int j = 0; j < P; ++j) { p[j] += 0.5 * eps * gr[j]; q[j] += eps * minv[j] * p[j]; }
    double lp = f(q, gr.data());
    for (int j = 0; j < P; ++j) p[j] += 0.5 * eps * gr[j];
    return lp;
  }
  // base_hmc::init_stepsize; q is restored on exit.  lp0 / g0 = density and gradient at q.
  void init_stepsize(double* q, const double* minv, double* eps_io, double lp0, const std::vector<double>& g0) {
    double eps = *eps_io;
    if (eps == 0 || eps > 1e7) return;
    const std::vector<double> q0(q, q + P);
    const double log08 = std::log(0.8);
    auto trial = [&]() {
      std::memcpy(q, q0.data(), sizeof(double) * P);
      sample_p(minv);
      gr = g0;
      const double H0 = -lp0 + kinetic(minv);
      const double lp = evolve(q, minv, eps);
      double h = -lp + kinetic(minv);
      if (std::isnan(h)) h = std::numeric_limits<double>::infinity();
      return H0 - h;
    };
    const int direction = trial() > log08 ? 1 : -1;
    for (;;) {
      const double dH = trial();
      if (direction == 1 && !(dH > log08)) break;
      if (direction == -1 && !(dH < log08)) break;
      eps = direction == 1 ? 2.0 * eps : 0.5 * eps;
      if (eps > 1e7 || eps == 0) break;  // the reference throws here ("Posterior is improper" / "No acceptably small step size")
    }
    std::memcpy(q, q0.data(), sizeof(double) * P);
    *eps_io = eps;
  }
};
}  // namespace

// learn_stepsize + learn_variance + (on a window end) init_stepsize and the dual-averaging restart:
// adapt_diag_e_static_hmc.hpp:33-47 == adapt_diag_e_nuts.hpp:31-44 (the static engine adds update_L_, derived from eps)
static void adapt_after_transition(Hmc& h, const LogpGrad& f, int P, double* q, double* ad, double* minv, double* wm,
                                   double* wm2, const double* cfg, const Windows& w, double accept_prob) {
  std::vector<double> g0(P);
  if (ad[9] == 1) {
    // learn_stepsize
    ad[4] += 1;
    const double eta = 1.0 / (ad[4] + cfg[6]);
    ad[2] = (1.0 - eta) * ad[2] + eta * (cfg[3] - accept_prob);
    const double x = ad[1] - ad[2] * std::sqrt(ad[4]) / cfg[4];
    const double x_eta = std::pow(ad[4], -cfg[5]);
    ad[3] = (1.0 - x_eta) * ad[3] + x_eta * x;
    ad[0] = std::exp(x);
    // learn_variance
    bool update = false;
    const bool in_window = ad[5] >= w.init_buffer && ad[5] < w.num_warmup - w.term_buffer && ad[5] != w.num_warmup;
    if (in_window) {
      ad[8] += 1;
      for (int j = 0; j < P; ++j) { const double d = q[j] - wm[j]; wm[j] += d / ad[8]; wm2[j] += d * (q[j] - wm[j]); }
    }
    if (ad[5] == ad[6] && ad[5] != w.num_warmup) {
      // compute_next_window
      const double last = w.num_warmup - w.term_buffer - 1;
      if (ad[6] != last) {
        ad[7] *= 2;
        ad[6] = ad[5] + ad[7];
        if (ad[6] != last) {
          const double boundary = ad[6] + 2 * ad[7];
          if (boundary >= w.num_warmup - w.term_buffer) ad[6] = last;
        }
      }
      const double n = ad[8];
      for (int j = 0; j < P; ++j) {
        double var = minv[j];
        if (n > 1) var = wm2[j] / (n - 1.0);
        minv[j] = (n / (n + 5.0)) * var + 1e-3 * (5.0 / (n + 5.0));
        wm[j] = 0; wm2[j] = 0;
      }
      ad[8] = 0;
      update = true;
    }
    ad[5] += 1;
    if (update) {
      const double lpq = f(q, g0.data());
      h.init_stepsize(q, minv, &ad[0], lpq, g0);
      ad[1] = std::log(10 * ad[0]);
      ad[2] = ad[3] = ad[4] = 0;
    }
  }
}

#define ORC_MAX_LEAPS 1000000

// One adaptive transition of one chain; returns 1 if the proposal was accepted.  out3 (optional): accept_stat, eps used, L used.
static int hmc_adapt_step(int P, const LogpGrad& f, double* q, double* ad, double* minv, double* wm, double* wm2,
                          const double* cfg, uint64_t seed, uint64_t chain, uint64_t step, double* lp_out, double* out3) {
  orc_rng g; orc_rng_init(&g, seed, chain, step);
  Hmc h(P, f, &g);
  const Windows w = windows_of(cfg);
  const double num_warmup = std::floor(cfg[2]);
  std::vector<double> g0(P);
  if (ad[9] == 0) {  // hmc_static_diag_e_adapt.hpp:62-74 + run_adaptive_sampler.hpp:55-57
    for (int j = 0; j < P; ++j) { minv[j] = 1.0; wm[j] = 0; wm2[j] = 0; }
    ad[0] = cfg[0]; ad[1] = std::log(10 * cfg[0]); ad[2] = ad[3] = ad[4] = 0;
    ad[5] = 0; ad[7] = w.windows_on ? w.base : 0; ad[6] = w.windows_on ? w.init_buffer + w.base - 1 : -1; ad[8] = 0;
    const double lp0 = f(q, g0.data());
    h.init_stepsize(q, minv, &ad[0], lp0, g0);
    ad[9] = 1;
  }
  if (ad[9] == 1 && (double)step >= num_warmup) { ad[0] = std::exp(ad[3]); ad[9] = 2; }  // complete_adaptation
  // ---- transition (base_static_hmc.hpp:42-72) ----
  const double eps = ad[0];
  double Ld = std::floor(cfg[1] / eps);  // update_L_
  if (!(Ld >= 1)) Ld = 1;
  if (Ld > ORC_MAX_LEAPS) Ld = ORC_MAX_LEAPS;
  const int L = (int)Ld;
  h.sample_p(minv);
  double lp = f(q, h.gr.data());
  const double lp0 = lp;
  const std::vector<double> q0(q, q + P);
  const double H0 = -lp + h.kinetic(minv);
  for (int l = 0; l < L; ++l) lp = h.evolve(q, minv, eps);
  double H = -lp + h.kinetic(minv);
  if (std::isnan(H)) H = std::numeric_limits<double>::infinity();
  double accept_prob = std::exp(H0 - H);
  const double u = orc_u01(&g);
  int acc = 1;
  if (accept_prob < 1 && u > accept_prob) { std::memcpy(q, q0.data(), sizeof(double) * P); lp = lp0; acc = 0; }
  if (accept_prob > 1) accept_prob = 1;
  if (out3) { out3[0] = accept_prob; out3[1] = eps; out3[2] = L; }
  adapt_after_transition(h, f, P, q, ad, minv, wm, wm2, cfg, w, accept_prob);
  if (lp_out) *lp_out = lp;
  return acc;
}

extern "C" int orc_linreg_hmc_adapt_step(const double* X, const double* y, int n, int d, double* q, double* ad,
                                         double* minv, double* wm, double* wm2, const double* cfg, uint64_t seed,
                                         uint64_t chain, uint64_t step, double* lp_out, double* out3) {
  LogpGrad f = [&](const double* qq, double* gr) { return orc_linreg_logp(X, y, n, d, qq, gr); };
  return hmc_adapt_step(d + 2, f, q, ad, minv, wm, wm2, cfg, seed, chain, step, lp_out, out3);
}

extern "C" int orc_birats_hmc_adapt_step(const double* x, const double* y, int N, int Tn, double* q, double* ad,
                                         double* minv, double* wm, double* wm2, const double* cfg, uint64_t seed,
                                         uint64_t chain, uint64_t step, double* lp_out, double* out3) {
  LogpGrad f = [&](const double* qq, double* gr) { return orc_birats_logp(x, y, N, Tn, qq, gr); };
  return hmc_adapt_step(2 * N + 6, f, q, ad, minv, wm, wm2, cfg, seed, chain, step, lp_out, out3);
}

// =====================================================================================================================
// NUTS with a diagonal metric and adaptation — `hmcStan`'s DEFAULT sampler (Stan.hs:428-440: algorithm=hmc engine=nuts,
// max_depth 10, metric diag_e, adapt engaged): cmdstan/stan/src/stan/mcmc/hmc/nuts/base_nuts.hpp:71-244 (transition,
// build_tree, compute_criterion — the 2.20 form: multinomial sampling, generalised U-turn criterion on rho, no extra
// sub-tree checks), adapt_diag_e_nuts.hpp:27-48.  Recursive exactly as the reference; the CUDA kernel builds the same
// tree iteratively and must reproduce it leaf for leaf (same uniforms in the same post-order).
// =====================================================================================================================
namespace {
double log_sum_exp2(double a, double b) {  // stan/math/prim/scal/fun/log_sum_exp.hpp
  if (a == -std::numeric_limits<double>::infinity()) return b;
  if (a == std::numeric_limits<double>::infinity() && b == std::numeric_limits<double>::infinity()) return a;
  if (a > b) return a + std::log1p(std::exp(b - a));
  return b + std::log1p(std::exp(a - b));
}

struct PsPoint { std::vector<double> q, p, g; double lp; };

struct Nuts {
  int P; const LogpGrad& f; orc_rng* rng; const double* minv; double eps; double max_deltaH;
  PsPoint z;
  int n_leapfrog; bool divergent; double sum_metro_prob;
  double H() const { double k = 0; for (int j = 0; j < P; ++j) k += minv[j] * z.p[j] * z.p[j]; return -z.lp + 0.5 * k; }
  void evolve(double e) {
    for (int j = 0; j < P; ++j) { z.p[j] += 0.5 * e * z.g[j]; z.q[j] += e * minv[j] * z.p[j]; }
    z.lp = f(z.q.data(), z.g.data());
    for (int j = 0; j < P; ++j) z.p[j] += 0.5 * e * z.g[j];
  }
  std::vector<double> p_sharp() const { std::vector<double> s(P); for (int j = 0; j < P; ++j) s[j] = minv[j] * z.p[j]; return s; }
  static bool criterion(const std::vector<double>& sm, const std::vector<double>& sp, const std::vector<double>& rho) {
    double a = 0, b = 0;
    for (size_t j = 0; j < rho.size(); ++j) { a += sp[j] * rho[j]; b += sm[j] * rho[j]; }
    return a > 0 && b > 0;
  }
  bool build_tree(int depth, PsPoint& z_propose, std::vector<double>& ps_left, std::vector<double>& ps_right,
                  std::vector<double>& rho, double H0, double sign, double& log_sum_weight) {
    if (depth == 0) {
      evolve(sign * eps);
      ++n_leapfrog;
      double h = H();
      if (std::isnan(h)) h = std::numeric_limits<double>::infinity();
      if ((h - H0) > max_deltaH) divergent = true;
      log_sum_weight = log_sum_exp2(log_sum_weight, H0 - h);
      if (H0 - h > 0) sum_metro_prob += 1; else sum_metro_prob += std::exp(H0 - h);
      z_propose = z;
      for (int j = 0; j < P; ++j) rho[j] += z.p[j];
      ps_left = p_sharp(); ps_right = ps_left;
      return !divergent;
    }
    std::vector<double> dummy(P);
    double lsw_left = -std::numeric_limits<double>::infinity();
    std::vector<double> rho_left(P, 0.0);
    if (!build_tree(depth - 1, z_propose, ps_left, dummy, rho_left, H0, sign, lsw_left)) return false;
    PsPoint z_propose_right = z;
    double lsw_right = -std::numeric_limits<double>::infinity();
    std::vector<double> rho_right(P, 0.0);
    if (!build_tree(depth - 1, z_propose_right, dummy, ps_right, rho_right, H0, sign, lsw_right)) return false;
    const double lsw_subtree = log_sum_exp2(lsw_left, lsw_right);
    log_sum_weight = log_sum_exp2(log_sum_weight, lsw_subtree);
    if (lsw_right > lsw_subtree) z_propose = z_propose_right;
    else if (orc_u01(rng) < std::exp(lsw_right - lsw_subtree)) z_propose = z_propose_right;
    std::vector<double> rho_subtree(P);
    for (int j = 0; j < P; ++j) { rho_subtree[j] = rho_left[j] + rho_right[j]; rho[j] += rho_subtree[j]; }
    return criterion(ps_left, ps_right, rho_subtree);
  }
};
}  // namespace

// One adaptive NUTS transition of one chain.  out5 (optional): accept_stat, eps, treedepth, n_leapfrog, divergent.
static void nuts_adapt_step(int P, const LogpGrad& f, double* q, double* ad, double* minv, double* wm, double* wm2,
                            const double* cfg, int max_depth, uint64_t seed, uint64_t chain, uint64_t step,
                            double* lp_out, double* out5) {
  orc_rng g; orc_rng_init(&g, seed, chain, step);
  Hmc h(P, f, &g);
  const Windows w = windows_of(cfg);
  const double num_warmup = std::floor(cfg[2]);
  std::vector<double> g0(P);
  if (ad[9] == 0) {  // services/sample/hmc_nuts_diag_e_adapt.hpp + run_adaptive_sampler.hpp:55-57
    for (int j = 0; j < P; ++j) { minv[j] = 1.0; wm[j] = 0; wm2[j] = 0; }
    ad[0] = cfg[0]; ad[1] = std::log(10 * cfg[0]); ad[2] = ad[3] = ad[4] = 0;
    ad[5] = 0; ad[7] = w.windows_on ? w.base : 0; ad[6] = w.windows_on ? w.init_buffer + w.base - 1 : -1; ad[8] = 0;
    const double lp0 = f(q, g0.data());
    h.init_stepsize(q, minv, &ad[0], lp0, g0);
    ad[9] = 1;
  }
  if (ad[9] == 1 && (double)step >= num_warmup) { ad[0] = std::exp(ad[3]); ad[9] = 2; }
  const double inf = std::numeric_limits<double>::infinity();
  Nuts nt{P, f, &g, minv, ad[0], 1000.0, PsPoint(), 0, false, 0.0};
  h.sample_p(minv);
  nt.z.q.assign(q, q + P); nt.z.p.assign(h.p.begin(), h.p.begin() + P); nt.z.g.resize(P);
  nt.z.lp = f(nt.z.q.data(), nt.z.g.data());
  PsPoint z_plus = nt.z, z_minus = nt.z, z_sample = nt.z, z_propose = nt.z;
  std::vector<double> ps_plus = nt.p_sharp(), ps_dummy = ps_plus, ps_minus = ps_plus, rho = nt.z.p;
  double log_sum_weight = 0;
  const double H0 = nt.H();
  int depth = 0;
  while (depth < max_depth) {
    std::vector<double> rho_subtree(P, 0.0);
    bool valid;
    double lsw_subtree = -inf;
    if (orc_u01(&g) > 0.5) {
      nt.z = z_plus;
      valid = nt.build_tree(depth, z_propose, ps_dummy, ps_plus, rho_subtree, H0, 1, lsw_subtree);
      z_plus = nt.z;
    } else {
      nt.z = z_minus;
      valid = nt.build_tree(depth, z_propose, ps_dummy, ps_minus, rho_subtree, H0, -1, lsw_subtree);
      z_minus = nt.z;
    }
    if (!valid) break;
    ++depth;
    if (lsw_subtree > log_sum_weight) z_sample = z_propose;
    else if (orc_u01(&g) < std::exp(lsw_subtree - log_sum_weight)) z_sample = z_propose;
    log_sum_weight = log_sum_exp2(log_sum_weight, lsw_subtree);
    for (int j = 0; j < P; ++j) rho[j] += rho_subtree[j];
    if (!Nuts::criterion(ps_minus, ps_plus, rho)) break;
  }
  const double accept_prob = nt.sum_metro_prob / (double)nt.n_leapfrog;
  std::memcpy(q, z_sample.q.data(), sizeof(double) * P);
  if (out5) { out5[0] = accept_prob; out5[1] = ad[0]; out5[2] = depth; out5[3] = nt.n_leapfrog; out5[4] = nt.divergent; }
  if (lp_out) *lp_out = z_sample.lp;
  adapt_after_transition(h, f, P, q, ad, minv, wm, wm2, cfg, w, accept_prob);
}

extern "C" void orc_linreg_nuts_step(const double* X, const double* y, int n, int d, double* q, double* ad, double* minv,
                                     double* wm, double* wm2, const double* cfg, int max_depth, uint64_t seed,
                                     uint64_t chain, uint64_t step, double* lp_out, double* out5) {
  LogpGrad f = [&](const double* qq, double* gr) { return orc_linreg_logp(X, y, n, d, qq, gr); };
  nuts_adapt_step(d + 2, f, q, ad, minv, wm, wm2, cfg, max_depth, seed, chain, step, lp_out, out5);
}

extern "C" void orc_birats_nuts_step(const double* x, const double* y, int N, int Tn, double* q, double* ad, double* minv,
                                     double* wm, double* wm2, const double* cfg, int max_depth, uint64_t seed,
                                     uint64_t chain, uint64_t step, double* lp_out, double* out5) {
  LogpGrad f = [&](const double* qq, double* gr) { return orc_birats_logp(x, y, N, Tn, qq, gr); };
  nuts_adapt_step(2 * N + 6, f, q, ad, minv, wm, wm2, cfg, max_depth, seed, chain, step, lp_out, out5);
}

// CPU-baseline driver (bench.py cpu_baseline / --impl reference): C chains x T NUTS transitions, one sequential chain at
// a time per host thread.  q, minv, wm, wm2 are [P][C], ad is [ORC_AD_LEN][C]; returns the total number of leapfrogs.
#include <atomic>
#include <thread>
template <class StepFn>
static long long nuts_run(int P, StepFn step_fn, double* q, double* ad, double* minv, double* wm, double* wm2, long long C,
                          long long T, uint64_t chain0, uint64_t step0, int nthreads) {
  std::atomic<long long> next{0}, leaps{0};
  auto work = [&]() {
    std::vector<double> qc(P), mc(P), w1(P), w2(P), ac(ORC_AD_LEN);
    double o5[5];
    for (;;) {
      const long long c = next.fetch_add(1);
      if (c >= C) break;
      for (int j = 0; j < P; ++j) { qc[j] = q[(size_t)j * C + c]; mc[j] = minv[(size_t)j * C + c]; w1[j] = wm[(size_t)j * C + c]; w2[j] = wm2[(size_t)j * C + c]; }
      for (int j = 0; j < ORC_AD_LEN; ++j) ac[j] = ad[(size_t)j * C + c];
      long long nl = 0;
      for (long long t = 0; t < T; ++t) {
        step_fn(qc.data(), ac.data(), mc.data(), w1.data(), w2.data(), chain0 + c, step0 + t, o5);
        nl += (long long)o5[3];
      }
      for (int j = 0; j < P; ++j) { q[(size_t)j * C + c] = qc[j]; minv[(size_t)j * C + c] = mc[j]; wm[(size_t)j * C + c] = w1[j]; wm2[(size_t)j * C + c] = w2[j]; }
      for (int j = 0; j < ORC_AD_LEN; ++j) ad[(size_t)j * C + c] = ac[j];
      leaps.fetch_add(nl);
    }
  };
  if (nthreads <= 1) { work(); return leaps.load(); }
  std::vector<std::thread> th;
  for (int i = 0; i < nthreads; ++i) th.emplace_back(work);
  for (auto& t : th) t.join();
  return leaps.load();
}

extern "C" long long orc_linreg_nuts_run(const double* X, const double* y, int n, int d, double* q, double* ad, double* minv,
                                         double* wm, double* wm2, long long C, long long T, const double* cfg,
                                         int max_depth, uint64_t seed, uint64_t chain0, uint64_t step0, int nthreads) {
  auto f = [&](double* qc, double* ac, double* mc, double* w1, double* w2, uint64_t chain, uint64_t step, double* o5) {
    orc_linreg_nuts_step(X, y, n, d, qc, ac, mc, w1, w2, cfg, max_depth, seed, chain, step, nullptr, o5);
  };
  return nuts_run(d + 2, f, q, ad, minv, wm, wm2, C, T, chain0, step0, nthreads);
}

extern "C" long long orc_birats_nuts_run(const double* x, const double* y, int N, int Tn, double* q, double* ad, double* minv,
                                         double* wm, double* wm2, long long C, long long T, const double* cfg,
                                         int max_depth, uint64_t seed, uint64_t chain0, uint64_t step0, int nthreads) {
  auto f = [&](double* qc, double* ac, double* mc, double* w1, double* w2, uint64_t chain, uint64_t step, double* o5) {
    orc_birats_nuts_step(x, y, N, Tn, qc, ac, mc, w1, w2, cfg, max_depth, seed, chain, step, nullptr, o5);
  };
  return nuts_run(2 * N + 6, f, q, ad, minv, wm, wm2, C, T, chain0, step0, nthreads);
}
