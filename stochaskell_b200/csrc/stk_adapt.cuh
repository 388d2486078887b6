// stk_adapt.cuh — per-chain warm-up adaptation for the HMC kernels: Stan's dual-averaging step size and windowed
// diagonal-metric estimation, the behaviour `hmcStan` gets from `adapt engaged=1` (Stan.hs:428-440), in the form of
// adapt_diag_e_static_hmc (cmdstan/stan/src/stan/mcmc/hmc/static/adapt_diag_e_static_hmc.hpp:29-51):
//   stepsize_adaptation.hpp:67-87 (learn_stepsize / complete_adaptation), windowed_adaptation.hpp:27-118 (window
//   schedule), var_adaptation.hpp:20-40 (Welford estimate shrunk towards 1e-3), base_hmc.hpp:73-126 (init_stepsize —
//   the model kernel runs its trials, this header holds the decision logic).
// Every chain adapts on its own, exactly as one CmdStan process per chain would; the state lives in the module's
// `ad` field (STK_AD_LEN doubles per chain) plus three P-vectors (minv, wm, wm2).
//
// ad:  0 eps (nominal step size)  1 mu  2 s_bar  3 x_bar  4 counter  5 window counter  6 next window end
//      7 window size  8 Welford n  9 phase (0 = not initialised, 1 = adapting, 2 = adaptation complete)
// The ten tunables are consecutive module params starting at `stepsize`:
//      stepsize int_time num_warmup delta gamma kappa t0 init_buffer term_buffer window
#ifndef STK_ADAPT_CUH
#define STK_ADAPT_CUH
#include "stk_device.cuh"

namespace stk {

#define STK_AD_LEN 10
#define STK_ADAPT_PARAMS \
  "param stepsize 1\nparam int_time 6.283185307179586\nparam num_warmup 1000\nparam delta 0.8\nparam gamma 0.05\n" \
  "param kappa 0.75\nparam t0 10\nparam init_buffer 75\nparam term_buffer 50\nparam window 25\n"
#define STK_MAX_LEAPS 1000000.0

struct AdaptCfg {
  double stepsize, int_time, num_warmup, delta, gamma, kappa, t0;
  double w_warmup, w_init, w_term, w_base;  // window schedule actually in force
  bool windows_on;
  // windowed_adaptation::set_window_params, including its two quirks: num_warmup < 20 -> no estimation at all; stages
  // that do not fit -> 15/75/10 % split stored WITHOUT restart(), so no window ever ends (metric never updated).
  STK_DEV void init(const double* par) {
    stepsize = par[0]; int_time = par[1]; num_warmup = floor(par[2]);
    delta = par[3]; gamma = par[4]; kappa = par[5]; t0 = par[6];
    const double ib = floor(par[7]), tb = floor(par[8]), bw = floor(par[9]);
    if (num_warmup < 20.0) { w_warmup = w_init = w_term = w_base = 0.0; windows_on = false; }
    else if (ib + bw + tb > num_warmup) {
      w_warmup = num_warmup; w_init = floor(0.15 * num_warmup); w_term = floor(0.10 * num_warmup);
      w_base = num_warmup - (w_init + w_term); windows_on = false;
    } else { w_warmup = num_warmup; w_init = ib; w_term = tb; w_base = bw; windows_on = true; }
  }
};

// hmc_static_diag_e_adapt.hpp:62-74: mu = log(10 * stepsize) is set from the configured step size, before init_stepsize
STK_DEV void adapt_reset(double (&ad)[STK_AD_LEN], const AdaptCfg& c) {
  ad[0] = c.stepsize; ad[1] = log(10.0 * c.stepsize); ad[2] = 0.0; ad[3] = 0.0; ad[4] = 0.0;
  ad[5] = 0.0; ad[7] = c.windows_on ? c.w_base : 0.0; ad[6] = c.windows_on ? c.w_init + c.w_base - 1.0 : -1.0;
  ad[8] = 0.0; ad[9] = 1.0;
}

// base_static_hmc::update_L_ (bounded: the reference would loop for ever on a vanishing step size)
STK_DEV int adapt_leaps(double int_time, double eps) {
  double L = floor(int_time / eps);
  if (!(L >= 1.0)) L = 1.0;
  if (L > STK_MAX_LEAPS) L = STK_MAX_LEAPS;
  return (int)L;
}

STK_DEV void adapt_learn_stepsize(double (&ad)[STK_AD_LEN], const AdaptCfg& c, double accept_stat) {
  ad[4] += 1.0;
  const double eta = 1.0 / (ad[4] + c.t0);
  ad[2] = (1.0 - eta) * ad[2] + eta * (c.delta - accept_stat);
  const double x = ad[1] - ad[2] * sqrt(ad[4]) / c.gamma;
  const double x_eta = pow(ad[4], -c.kappa);
  ad[3] = (1.0 - x_eta) * ad[3] + x_eta * x;
  ad[0] = exp(x);
}

STK_DEV bool adapt_in_window(const double (&ad)[STK_AD_LEN], const AdaptCfg& c) {
  return ad[5] >= c.w_init && ad[5] < c.w_warmup - c.w_term && ad[5] != c.w_warmup;
}
STK_DEV bool adapt_window_ends(const double (&ad)[STK_AD_LEN], const AdaptCfg& c) {
  return ad[5] == ad[6] && ad[5] != c.w_warmup;
}
// windowed_adaptation::compute_next_window
STK_DEV void adapt_next_window(double (&ad)[STK_AD_LEN], const AdaptCfg& c) {
  const double last = c.w_warmup - c.w_term - 1.0;
  if (ad[6] == last) return;
  ad[7] *= 2.0;
  ad[6] = ad[5] + ad[7];
  if (ad[6] == last) return;
  if (ad[6] + 2.0 * ad[7] >= c.w_warmup - c.w_term) ad[6] = last;
}
// one coordinate of welford_var_estimator::add_sample (n already incremented)
STK_DEV void welford_add(double n, double q, double& m, double& m2) {
  const double d = q - m; m += d / n; m2 += d * (q - m);
}
// one coordinate of var_adaptation::learn_variance at a window end
STK_DEV double adapt_window_var(double n, double m2, double old_minv) {
  const double var = n > 1.0 ? m2 / (n - 1.0) : old_minv;
  return (n / (n + 5.0)) * var + 1e-3 * (5.0 / (n + 5.0));
}

// Decision logic of base_hmc::init_stepsize, one call per trial (the model kernel computes dH = H0 - h of a single
// leapfrog step with fresh momenta).  dir = 0 before the first trial.  Returns true when the search is finished.
STK_DEV bool init_stepsize_update(double& eps, int& dir, double dH) {
  const double log08 = -0.2231435513142097557662950903098345;  // log(0.8)
  if (dir == 0) { dir = dH > log08 ? 1 : -1; return false; }
  if (dir == 1 && !(dH > log08)) return true;
  if (dir == -1 && !(dH < log08)) return true;
  eps = dir == 1 ? 2.0 * eps : 0.5 * eps;
  return eps > 1e7 || eps == 0.0;
}

}  // namespace stk
#endif
