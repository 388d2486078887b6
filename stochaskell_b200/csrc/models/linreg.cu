// linreg.cu — hand-tuned module for config C1: Bayesian linear regression
//   b ~ normal(0,10); w_j ~ normal(0,10); sigma ~ truncated 0 inf (cauchy 0 5)
//   y_i ~ normal(b + x_i . w, sigma)
// the d-dimensional form of linreg.ipynb cell 1 (the Stan text it generates is
// linreg.ipynb cell 7; oracle/stan/linreg.stan is that text with vector w).
//
// Pattern ("HMC over a sum-over-data likelihood"):
//   * one CTA = 32 chains (lanes) x K observation slices (warps);
//   * X|y staged once per launch into shared memory as rows of D+2 doubles,
//     every lane of a warp reads the SAME row -> LDS.128 broadcasts, no bank
//     conflicts, X never re-read from HBM during the launch;
//   * each thread accumulates its slice's partial sums for ITS chain in
//     registers (no shuffle per observation); the K slices are combined once
//     per gradient evaluation through shared memory in a fixed order
//     (deterministic);
//   * the whole transition loop (momentum draw -> L leapfrogs -> accept) runs
//     inside the kernel; chain state lives in registers for the launch.
// HMC semantics: cmdstan/stan/src/stan/mcmc/hmc/static/base_static_hmc.hpp:42-72,
// integrators/base_leapfrog.hpp:18-28 + expl_leapfrog.hpp:17-34, unit_e metric.
//
// Unconstrained parameters q = (b, w[0..D-1], log sigma), P = D+2.
#include "../stk_device.cuh"
#include "../stk_adapt.cuh"

using namespace stk;

#ifndef LINREG_D
#define LINREG_D 10
#endif

namespace {

constexpr int D = LINREG_D;
constexpr int P = D + 2;        // b, w[D], log sigma
constexpr int RS = (D + 2) & ~1;  // smem row stride in doubles (x[D], y, pad) -> even
static_assert(RS >= D + 1, "row stride");
constexpr int NACC = D + 2;     // sum r, sum x_j r (D), sum r^2

enum { F_Q = 0, F_LP = 1, F_X = 2, F_Y = 3, F_YSIM = 4, F_AD = 5, F_MINV = 6, F_WM = 7, F_WM2 = 8 };
enum { PAR_EPS = 0, PAR_L = 1, PAR_RW = 2, PAR_ADAPT = 3 /* stepsize .. window, stk_adapt.cuh */ };

// stage [n][D] X and [n] y from HBM into smem rows
__device__ __forceinline__ void stage_data(const StkArgs& a, double* xy, int n) {
  const double* X = (const double*)a.field[F_X];
  const double* y = (const double*)a.field[F_Y];
  for (int idx = threadIdx.x; idx < n * D; idx += blockDim.x) {
    int i = idx / D, j = idx - i * D;
    xy[i * RS + j] = X[idx];
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    xy[i * RS + D] = y[i];
    if (RS > D + 1) xy[i * RS + D + 1] = 0.0;
  }
}

// Partial sums over this thread's observation slice, then the fixed-order
// cross-slice combine.  GRAD=false skips the D+1 gradient accumulators.
template <bool GRAD>
__device__ __forceinline__ void data_pass(const double* __restrict__ xy, int n, int slice, int K,
                                          int lane, double* red, double* tot,
                                          const double (&q)[P], double (&acc)[NACC]) {
  double b = q[0];
  double w[D];
#pragma unroll
  for (int j = 0; j < D; ++j) w[j] = q[1 + j];
  double sr = 0.0, ss = 0.0;
  double gw[D];
#pragma unroll
  for (int j = 0; j < D; ++j) gw[j] = 0.0;

#pragma unroll 2
  for (int i = slice; i < n; i += K) {
    const double2* row = reinterpret_cast<const double2*>(xy + i * RS);
    double xr[RS];
#pragma unroll
    for (int j = 0; j < RS / 2; ++j) { double2 v = row[j]; xr[2 * j] = v.x; xr[2 * j + 1] = v.y; }
    // two interleaved partial dot products shorten the dependent DFMA chain
    double m0 = b, m1 = 0.0;
#pragma unroll
    for (int j = 0; j < D; j += 2) {
      m0 = fma(xr[j], w[j], m0);
      if (j + 1 < D) m1 = fma(xr[j + 1], w[j + 1], m1);
    }
    double r = xr[D] - (m0 + m1);
    ss = fma(r, r, ss);
    if (GRAD) {
      sr += r;
#pragma unroll
      for (int j = 0; j < D; ++j) gw[j] = fma(xr[j], r, gw[j]);
    }
  }

  // ---- combine the K slices (fixed order k = 0..K-1) ----
  constexpr int NA = GRAD ? NACC : 1;
  if (GRAD) {
    red[(slice * NACC + 0) * 32 + lane] = sr;
#pragma unroll
    for (int j = 0; j < D; ++j) red[(slice * NACC + 1 + j) * 32 + lane] = gw[j];
    red[(slice * NACC + D + 1) * 32 + lane] = ss;
  } else {
    red[(slice * NACC + D + 1) * 32 + lane] = ss;
  }
  __syncthreads();
  for (int t = slice; t < NA; t += K) {
    int aidx = GRAD ? t : D + 1;
    double s = 0.0;
    for (int k = 0; k < K; ++k) s += red[(k * NACC + aidx) * 32 + lane];
    tot[aidx * 32 + lane] = s;
  }
  __syncthreads();
  if (GRAD) {
#pragma unroll
    for (int t = 0; t < NACC; ++t) acc[t] = tot[t * 32 + lane];
  } else {
    acc[D + 1] = tot[(D + 1) * 32 + lane];
  }
}

// joint log-density on the unconstrained scale (all constants kept, as
// Program.hs `lpdf` does) and its gradient.
template <bool GRAD>
__device__ __forceinline__ double finish_eval(int n, const double (&q)[P], const double (&acc)[NACC],
                                              double (&g)[P]) {
  const double ls = q[P - 1];
  const double sigma = exp(ls);
  const double is2 = exp(-2.0 * ls);
  const double ss = acc[D + 1];
  // likelihood: sum_i normal_lpdf(y_i | mu_i, sigma)
  double lp = -(double)n * (kHalfLog2Pi + ls) - 0.5 * ss * is2;
  // priors: b, w_j ~ normal(0,10)
  const double lps = -kHalfLog2Pi - 2.302585092994045684017991454684;  // -0.5log2pi - log 10
  double sq = q[0] * q[0];
#pragma unroll
  for (int j = 0; j < D; ++j) sq = fma(q[1 + j], q[1 + j], sq);
  lp += (double)(D + 1) * lps - 0.005 * sq;
  // sigma ~ cauchy(0,5) T[0,]: cauchy_lpdf - log(1 - cdf(0)) = cauchy_lpdf + log 2
  const double z = sigma * 0.2;
  lp += -kLogPi - 1.6094379124341003746007593332262 - log1p(z * z) + 0.69314718055994530941723212145818;
  // Jacobian of sigma = exp(ls)   (stan_math lb_constrain.hpp)
  lp += ls;
  if (GRAD) {
    g[0] = acc[0] * is2 - 0.01 * q[0];
#pragma unroll
    for (int j = 0; j < D; ++j) g[1 + j] = acc[1 + j] * is2 - 0.01 * q[1 + j];
    // d/dls: -n + ss/sigma^2 - (2 z^2/(1+z^2)) + 1
    g[P - 1] = -(double)n + ss * is2 - 2.0 * z * z / (1.0 + z * z) + 1.0;
  }
  return lp;
}

struct Smem {
  double* xy; double* red; double* tot;
  __device__ Smem(double* base, int n, int K) {
    xy = base; red = xy + (size_t)n * RS; tot = red + (size_t)K * NACC * 32;
  }
};

}  // namespace

// ---------------------------------------------------------------------------
// static HMC, n_steps transitions per launch
// ---------------------------------------------------------------------------
extern "C" __global__ void __launch_bounds__(384, 1)
linreg_hmc(const __grid_constant__ StkArgs a) {
  extern __shared__ __align__(16) double smem[];
  const int n = (int)a.flen[F_Y];
  const int K = blockDim.x >> 5;
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  Smem sm(smem, n, K);
  stage_data(a, sm.xy, n);
  __syncthreads();

  const long long chain = (long long)blockIdx.x * 32 + lane;
  const bool active = chain < a.n_chains;
  const long long cidx = active ? chain : a.n_chains - 1;  // idle lanes shadow the last chain
  double* Q = (double*)a.field[F_Q];
  double q[P], g[P], acc[NACC];
#pragma unroll
  for (int j = 0; j < P; ++j) q[j] = Q[(long long)j * a.n_chains + cidx];

  const double eps = a.par[PAR_EPS];
  const int L = (int)a.par[PAR_L];
  unsigned long long n_acc = 0;
  double lp = 0.0;

  for (long long t = 0; t < a.n_steps; ++t) {
    Rng rng; rng.init(a.seed, (u64)(a.chain0 + cidx), a.step0 + (u64)t);
    // hamiltonian_.sample_p : p ~ N(0, I)   (unit_e_metric.hpp:44-49)
    double p[P];
#pragma unroll
    for (int j = 0; j < P; j += 2) {
      double z0, z1; rng.normal2(z0, z1);
      p[j] = z0; if (j + 1 < P) p[j + 1] = z1;
    }
    // hamiltonian_.init : potential + gradient at the current point
    data_pass<true>(sm.xy, n, slice, K, lane, sm.red, sm.tot, q, acc);
    lp = finish_eval<true>(n, q, acc, g);
    double ke = 0.0;
#pragma unroll
    for (int j = 0; j < P; ++j) ke = fma(p[j], p[j], ke);
    const double H0 = -lp + 0.5 * ke;
    double q0[P]; const double lp0 = lp;
#pragma unroll
    for (int j = 0; j < P; ++j) q0[j] = q[j];

    for (int l = 0; l < L; ++l) {
#pragma unroll
      for (int j = 0; j < P; ++j) { p[j] = fma(0.5 * eps, g[j], p[j]); q[j] = fma(eps, p[j], q[j]); }
      data_pass<true>(sm.xy, n, slice, K, lane, sm.red, sm.tot, q, acc);
      lp = finish_eval<true>(n, q, acc, g);
#pragma unroll
      for (int j = 0; j < P; ++j) p[j] = fma(0.5 * eps, g[j], p[j]);
    }
    ke = 0.0;
#pragma unroll
    for (int j = 0; j < P; ++j) ke = fma(p[j], p[j], ke);
    double H = -lp + 0.5 * ke;
    if (isnan(H)) H = pos_inf();
    const double accept_prob = exp(H0 - H);
    const double u = rng.u01();
    if (accept_prob < 1.0 && u > accept_prob) {
#pragma unroll
      for (int j = 0; j < P; ++j) q[j] = q0[j];
      lp = lp0;
    } else {
      ++n_acc;
    }
    long long slot;
    if (slice == 0 && active && record_slot(a, t, &slot)) {
#pragma unroll
      for (int j = 0; j < P; ++j) record_value(a, slot, j, chain, q[j]);
    }
  }
  if (slice == 0 && active) {
#pragma unroll
    for (int j = 0; j < P; ++j) Q[(long long)j * a.n_chains + chain] = q[j];
    ((double*)a.field[F_LP])[chain] = lp;
    if (a.accept) a.accept[chain] += n_acc;
  }
}

// ---------------------------------------------------------------------------
// adaptive static HMC with a diagonal metric: Stan's adapt_diag_e_static_hmc, every chain adapting its own step
// size (dual averaging) and metric (windowed Welford variances) during the first num_warmup transitions
// (stk_adapt.cuh; oracle/stk_oracle_adapt.cpp is the step-by-step restatement).  The K observation slices of a
// chain hold replicas of the whole per-chain state and take every decision redundantly; anything that contains a
// data pass runs in CTA-uniform control flow (the loops run to the CTA's maximum trip count, lanes that are done
// keep evaluating at their final point and ignore the result).
// ---------------------------------------------------------------------------
extern "C" __global__ void __launch_bounds__(256, 1)
linreg_hmc_adapt(const __grid_constant__ StkArgs a) {
  extern __shared__ __align__(16) double smem[];
  const int n = (int)a.flen[F_Y];
  const int K = blockDim.x >> 5;
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  Smem sm(smem, n, K);
  stage_data(a, sm.xy, n);
  __syncthreads();

  const long long chain = (long long)blockIdx.x * 32 + lane;
  const bool active = chain < a.n_chains;
  const long long cidx = active ? chain : a.n_chains - 1;
  const long long nc = a.n_chains;
  const bool writer = slice == 0 && active;
  double* Q = (double*)a.field[F_Q];
  double* AD = (double*)a.field[F_AD];
  double* MINV = (double*)a.field[F_MINV];
  double* WM = (double*)a.field[F_WM];
  double* WM2 = (double*)a.field[F_WM2];
  AdaptCfg cfg; cfg.init(&a.par[PAR_ADAPT]);
  double q[P], g[P], p[P], minv[P], acc[NACC], ad[STK_AD_LEN];
#pragma unroll
  for (int j = 0; j < P; ++j) { q[j] = Q[j * nc + cidx]; minv[j] = MINV[j * nc + cidx]; }
#pragma unroll
  for (int j = 0; j < STK_AD_LEN; ++j) ad[j] = AD[j * nc + cidx];
  unsigned long long n_acc = 0;
  double lp = 0.0;

  auto sample_p = [&](Rng& rng) {
#pragma unroll
    for (int j = 0; j < P; j += 2) {
      double z0, z1; rng.normal2(z0, z1);
      p[j] = z0 / sqrt(minv[j]); if (j + 1 < P) p[j + 1] = z1 / sqrt(minv[j + 1]);
    }
  };
  auto kinetic = [&]() {
    double ke = 0.0;
#pragma unroll
    for (int j = 0; j < P; ++j) ke = fma(minv[j] * p[j], p[j], ke);
    return 0.5 * ke;
  };
  // base_hmc::init_stepsize for the lanes with `want`; lp / g hold the density and gradient at q on entry and on exit
  auto init_stepsize = [&](Rng& rng, bool want) {
    double eps = ad[0];
    int dir = 0;
    bool done = !want || eps == 0.0 || eps > 1e7;
    const double lp0 = lp;
    double g0[P];
#pragma unroll
    for (int j = 0; j < P; ++j) g0[j] = g[j];
    while (__syncthreads_or(!done)) {
      double qt[P], gt[P];
      if (!done) sample_p(rng);
      const double H0 = -lp0 + kinetic();
#pragma unroll
      for (int j = 0; j < P; ++j) { p[j] = fma(0.5 * eps, g0[j], p[j]); qt[j] = fma(eps * minv[j], p[j], q[j]); }
      data_pass<true>(sm.xy, n, slice, K, lane, sm.red, sm.tot, qt, acc);
      const double lpt = finish_eval<true>(n, qt, acc, gt);
#pragma unroll
      for (int j = 0; j < P; ++j) p[j] = fma(0.5 * eps, gt[j], p[j]);
      double h = -lpt + kinetic();
      if (isnan(h)) h = pos_inf();
      if (!done) done = init_stepsize_update(eps, dir, H0 - h);
    }
    if (want) ad[0] = eps;
  };

  for (long long t = 0; t < a.n_steps; ++t) {
    const unsigned long long step = a.step0 + (u64)t;
    Rng rng; rng.init(a.seed, (u64)(a.chain0 + cidx), step);
    const bool fresh = ad[9] == 0.0;
    if (fresh) {
#pragma unroll
      for (int j = 0; j < P; ++j) minv[j] = 1.0;
      adapt_reset(ad, cfg);
    }
    // hamiltonian_.init: density and gradient at the current point (shared by init_stepsize and the transition)
    data_pass<true>(sm.xy, n, slice, K, lane, sm.red, sm.tot, q, acc);
    lp = finish_eval<true>(n, q, acc, g);
    if (__syncthreads_or(fresh)) init_stepsize(rng, fresh);
    if (ad[9] == 1.0 && (double)step >= cfg.num_warmup) { ad[0] = exp(ad[3]); ad[9] = 2.0; }  // complete_adaptation
    const double eps = ad[0];
    const int L = adapt_leaps(cfg.int_time, eps);
    sample_p(rng);
    const double H0 = -lp + kinetic();
    double q0[P]; const double lp0 = lp;
#pragma unroll
    for (int j = 0; j < P; ++j) q0[j] = q[j];
    for (int l = 0; __syncthreads_or(l < L); ++l) {
      const bool on = l < L;
      if (on) {
#pragma unroll
        for (int j = 0; j < P; ++j) { p[j] = fma(0.5 * eps, g[j], p[j]); q[j] = fma(eps * minv[j], p[j], q[j]); }
      }
      data_pass<true>(sm.xy, n, slice, K, lane, sm.red, sm.tot, q, acc);
      lp = finish_eval<true>(n, q, acc, g);
      if (on) {
#pragma unroll
        for (int j = 0; j < P; ++j) p[j] = fma(0.5 * eps, g[j], p[j]);
      }
    }
    double H = -lp + kinetic();
    if (isnan(H)) H = pos_inf();
    double accept_prob = exp(H0 - H);
    const double u = rng.u01();
    if (accept_prob < 1.0 && u > accept_prob) {
#pragma unroll
      for (int j = 0; j < P; ++j) q[j] = q0[j];
      lp = lp0;
    } else {
      ++n_acc;
    }
    if (accept_prob > 1.0) accept_prob = 1.0;
    bool update = false, store = false;
    if (ad[9] == 1.0) {
      adapt_learn_stepsize(ad, cfg, accept_prob);
      const bool in_w = adapt_in_window(ad, cfg), ends = adapt_window_ends(ad, cfg);
      if (in_w || ends) {
        double wm[P], wm2[P];
#pragma unroll
        for (int j = 0; j < P; ++j) { wm[j] = WM[j * nc + cidx]; wm2[j] = WM2[j * nc + cidx]; }
        if (in_w) {
          ad[8] += 1.0;
#pragma unroll
          for (int j = 0; j < P; ++j) welford_add(ad[8], q[j], wm[j], wm2[j]);
        }
        if (ends) {
          adapt_next_window(ad, cfg);
#pragma unroll
          for (int j = 0; j < P; ++j) { minv[j] = adapt_window_var(ad[8], wm2[j], minv[j]); wm[j] = 0.0; wm2[j] = 0.0; }
          ad[8] = 0.0;
          update = true;
        }
        // every slice reads the estimator before slice 0 writes it back (after the barrier below)
#pragma unroll
        for (int j = 0; j < P; ++j) { g[j] = wm[j]; p[j] = wm2[j]; }  // park in dead registers until the barrier
      }
      ad[5] += 1.0;
      store = in_w || ends;
    }
    __syncthreads();
    if (store && writer) {
#pragma unroll
      for (int j = 0; j < P; ++j) { WM[j * nc + chain] = g[j]; WM2[j * nc + chain] = p[j]; }
    }
    if (__syncthreads_or(update)) {
      data_pass<true>(sm.xy, n, slice, K, lane, sm.red, sm.tot, q, acc);
      lp = finish_eval<true>(n, q, acc, g);
      init_stepsize(rng, update);
      if (update) { ad[1] = log(10.0 * ad[0]); ad[2] = 0.0; ad[3] = 0.0; ad[4] = 0.0; }
    }
    long long slot;
    if (writer && record_slot(a, t, &slot)) {
#pragma unroll
      for (int j = 0; j < P; ++j) record_value(a, slot, j, chain, q[j]);
    }
  }
  if (writer) {
#pragma unroll
    for (int j = 0; j < P; ++j) { Q[j * nc + chain] = q[j]; MINV[j * nc + chain] = minv[j]; }
#pragma unroll
    for (int j = 0; j < STK_AD_LEN; ++j) AD[j * nc + chain] = ad[j];
    ((double*)a.field[F_LP])[chain] = lp;
    if (a.accept) a.accept[chain] += n_acc;
  }
}

// ---------------------------------------------------------------------------
// random-walk Metropolis (the `mh`-style kernel of mh.ipynb cell 1, proposal
// q' = q + rw * N(0, I) on the unconstrained scale), n_steps per launch
// ---------------------------------------------------------------------------
extern "C" __global__ void __launch_bounds__(512, 1)
linreg_mh(const __grid_constant__ StkArgs a) {
  extern __shared__ __align__(16) double smem[];
  const int n = (int)a.flen[F_Y];
  const int K = blockDim.x >> 5;
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  Smem sm(smem, n, K);
  stage_data(a, sm.xy, n);
  __syncthreads();
  const long long chain = (long long)blockIdx.x * 32 + lane;
  const bool active = chain < a.n_chains;
  const long long cidx = active ? chain : a.n_chains - 1;
  double* Q = (double*)a.field[F_Q];
  double q[P], g[P], acc[NACC];
#pragma unroll
  for (int j = 0; j < P; ++j) q[j] = Q[(long long)j * a.n_chains + cidx];
  const double rw = a.par[PAR_RW];
  unsigned long long n_acc = 0;
  data_pass<false>(sm.xy, n, slice, K, lane, sm.red, sm.tot, q, acc);
  double lp = finish_eval<false>(n, q, acc, g);
  for (long long t = 0; t < a.n_steps; ++t) {
    Rng rng; rng.init(a.seed, (u64)(a.chain0 + cidx), a.step0 + (u64)t);
    double qn[P];
#pragma unroll
    for (int j = 0; j < P; j += 2) {
      double z0, z1; rng.normal2(z0, z1);
      qn[j] = fma(rw, z0, q[j]); if (j + 1 < P) qn[j + 1] = fma(rw, z1, q[j + 1]);
    }
    data_pass<false>(sm.xy, n, slice, K, lane, sm.red, sm.tot, qn, acc);
    const double lpn = finish_eval<false>(n, qn, acc, g);
    const double ratio = exp(lpn - lp);           // rjmc: bernoulli(min 1 a), Program.hs:1278-1283
    const double u = rng.u01();
    if (u < fmin(1.0, ratio)) {
#pragma unroll
      for (int j = 0; j < P; ++j) q[j] = qn[j];
      lp = lpn; ++n_acc;
    }
    long long slot;
    if (slice == 0 && active && record_slot(a, t, &slot)) {
#pragma unroll
      for (int j = 0; j < P; ++j) record_value(a, slot, j, chain, q[j]);
    }
  }
  if (slice == 0 && active) {
#pragma unroll
    for (int j = 0; j < P; ++j) Q[(long long)j * a.n_chains + chain] = q[j];
    ((double*)a.field[F_LP])[chain] = lp;
    if (a.accept) a.accept[chain] += n_acc;
  }
}

// ---------------------------------------------------------------------------
// logp / logp_grad at the current state (parity tier 1)
// ---------------------------------------------------------------------------
template <bool GRAD>
__device__ __forceinline__ void logp_entry(const StkArgs& a) {
  extern __shared__ __align__(16) double smem[];
  const int n = (int)a.flen[F_Y];
  const int K = blockDim.x >> 5;
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  Smem sm(smem, n, K);
  stage_data(a, sm.xy, n);
  __syncthreads();
  const long long chain = (long long)blockIdx.x * 32 + lane;
  const bool active = chain < a.n_chains;
  const long long cidx = active ? chain : a.n_chains - 1;
  const double* Q = (const double*)a.field[F_Q];
  double q[P], g[P], acc[NACC];
#pragma unroll
  for (int j = 0; j < P; ++j) q[j] = Q[(long long)j * a.n_chains + cidx];
  data_pass<GRAD>(sm.xy, n, slice, K, lane, sm.red, sm.tot, q, acc);
  const double lp = finish_eval<GRAD>(n, q, acc, g);
  if (slice == 0 && active) {
    a.out_lp[chain] = lp;
    if (GRAD) {
#pragma unroll
      for (int j = 0; j < P; ++j) a.out_grad[(long long)j * a.n_chains + chain] = g[j];
    }
  }
}
extern "C" __global__ void __launch_bounds__(512, 1) linreg_logp(const __grid_constant__ StkArgs a) {
  logp_entry<false>(a);
}
extern "C" __global__ void __launch_bounds__(512, 1) linreg_logp_grad(const __grid_constant__ StkArgs a) {
  logp_entry<true>(a);
}

// ---------------------------------------------------------------------------
// forward simulation (simulate, Program.hs:1038): b, w, sigma from the priors,
// y_i ~ normal(b + x_i.w, sigma) for the given X.  Stream: transition counter
// = 2^64-1 marks "simulate"; slot 0..: b,w pairs, then sigma attempts, then
// slot 64+i for y_i.
// ---------------------------------------------------------------------------
extern "C" __global__ void __launch_bounds__(512, 1)
linreg_simulate(const __grid_constant__ StkArgs a) {
  extern __shared__ __align__(16) double smem[];
  const int n = (int)a.flen[F_Y];
  const int K = blockDim.x >> 5;
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  Smem sm(smem, n, K);
  stage_data(a, sm.xy, n);
  __syncthreads();
  const long long chain = (long long)blockIdx.x * 32 + lane;
  const bool active = chain < a.n_chains;
  const long long cidx = active ? chain : a.n_chains - 1;
  Rng rng; rng.init(a.seed, (u64)(a.chain0 + cidx), ~0ULL);
  double q[P];
#pragma unroll
  for (int j = 0; j < D + 1; j += 2) {
    double z0, z1; rng.normal2(z0, z1);
    q[j] = 10.0 * z0; if (j + 1 < D + 1) q[j + 1] = 10.0 * z1;
  }
  // truncated 0 inf (cauchy 0 5): rejection as CC.hs:282-286 (do..while out of range)
  double sigma = 1.0;
  for (int it = 0; it < 64; ++it) { sigma = draw_cauchy(rng, 0.0, 5.0); if (sigma >= 0.0) break; }
  sigma = fabs(sigma);
  q[P - 1] = log(sigma);
  double* ysim = (double*)a.field[F_YSIM];
  for (int i = slice; i < n; i += K) {
    const double* row = sm.xy + i * RS;
    double mu = q[0];
#pragma unroll
    for (int j = 0; j < D; ++j) mu = fma(row[j], q[1 + j], mu);
    Rng r2 = rng; r2.slot = 64u + (u32)i;
    const double yi = mu + sigma * r2.normal();
    if (active && ysim) ysim[(long long)i * a.n_chains + chain] = yi;
  }
  if (slice == 0 && active) {
    double* Q = (double*)a.field[F_Q];
#pragma unroll
    for (int j = 0; j < P; ++j) Q[(long long)j * a.n_chains + chain] = q[j];
  }
}

// descriptor ------------------------------------------------------------------
// The numeric literals below must be plain numbers (the runtime's descriptor
// parser does not evaluate expressions); the build passes them with -D for
// other D, and the static_asserts keep them honest.
#ifndef LINREG_P
#define LINREG_P 12            /* D + 2 */
#define LINREG_BLOCK 256       /* K = 8 observation slices */
#define LINREG_ROW_BYTES 96    /* RS * 8 */
#define LINREG_SMEM_FIXED 27648 /* (K*NACC + NACC) * 32 * 8 */
#endif
static_assert(LINREG_P == P, "LINREG_P");
static_assert(LINREG_ROW_BYTES == RS * 8, "LINREG_ROW_BYTES");
static_assert(LINREG_SMEM_FIXED == ((LINREG_BLOCK / 32) * NACC + NACC) * 32 * 8, "LINREG_SMEM_FIXED");
#define STR2(x) #x
#define STR(x) STR2(x)
#define GEOM " " STR(LINREG_BLOCK) " 32 " STR(LINREG_SMEM_FIXED) " y " STR(LINREG_ROW_BYTES) "\n"
extern "C" __device__ const char stk_desc[] =
    "stk-module 1\n"
    "name linreg\n"
    "nparam " STR(LINREG_P) "\n"
    "field q state f64 " STR(LINREG_P) "\n"
    "field lp state f64 1\n"
    "field X data f64 0\n"
    "field y data f64 0\n"
    "field ysim state f64 @y\n"
    "field ad state f64 10\n"
    "field minv state f64 " STR(LINREG_P) "\n"
    "field wm state f64 " STR(LINREG_P) "\n"
    "field wm2 state f64 " STR(LINREG_P) "\n"
    "param eps 0.01\n"
    "param L 10\n"
    "param rw 0.05\n"
    STK_ADAPT_PARAMS
    "record q\n"
    "kernel hmc linreg_hmc" GEOM
    "kernel mh linreg_mh" GEOM
    "kernel hmc_adapt linreg_hmc_adapt" GEOM
    "entry logp linreg_logp" GEOM
    "entry logp_grad linreg_logp_grad" GEOM
    "entry simulate linreg_simulate" GEOM;
