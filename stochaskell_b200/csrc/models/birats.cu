// birats.cu — hand-tuned module for config C4: the hierarchical BiRats growth model
//   birats.ipynb cells 1-2 / app/benchmark/Main.hs:106-125, on the unconstrained scale of the Stan program that
//   Language.Stan generates for it (birats.ipynb cell 8):
//     q = (betaMu[2], tau[2], y_omega, beta[.,1] (N), beta[.,2] (N), log sigma),  P = 2N + 6  (66 for 30 rats)
//   lp: see oracle/stk_oracle_birats.cpp (pinned to the vendored CmdStan's log_prob_grad at 1e-10).
//   HMC: base_static_hmc.hpp:42-72, expl_leapfrog.hpp:17-34, unit_e metric — `hmcStan`'s static engine
//   (benchStanHMC, app/benchmark/Main.hs:250-258).
//
// Pattern ("hierarchical model, one WARP per chain"):
//   * lane i owns rat i: its two coefficients, their momenta and gradient, and its row of y in registers;
//   * the 6 hyper-parameters (+ momenta, gradients) are replicated in every lane and updated redundantly — the
//     xor-butterfly sums are bitwise identical on all lanes, so the replicas never diverge;
//   * one gradient evaluation = 5 observations + one 2-D Gaussian term per lane, then 6 warp-shuffle sums
//     (d/d betaMu, d/d tau, d/d z, sum r^2) — 7 with lp, which only the ends of a trajectory need;
//   * momenta: lane l draws Philox block l (two normals), p[j] is fetched from lane j/2 by shuffle — the same
//     assignment of normals to parameters as the oracle's sequential loop;
//   * x and y are staged in shared memory once per launch and then held in registers.
#include "../stk_device.cuh"
#include "../stk_adapt.cuh"

using namespace stk;

namespace {

constexpr int TMAX = 8;                 // observations per rat held in registers
enum { F_Q = 0, F_LP = 1, F_X = 2, F_Y = 3, F_AD = 4, F_MINV = 5, F_WM = 6, F_WM2 = 7 };
enum { PAR_EPS = 0, PAR_L = 1, PAR_ADAPT = 2 /* stepsize .. window, stk_adapt.cuh */ };

struct Hyper { double mu1, mu2, t1, t2, yo, ls; };

template <int G>
struct Rat {
  double b1, b2;                        // this lane's coefficients
  double y[TMAX];
};

__device__ __forceinline__ double warp_sum(double v) { return group_sum<32>(v); }

// log-density (optional) and gradient at (h, b1, b2).  gh[6] = d/d(mu1, mu2, t1, t2, yo, ls); g1, g2 = d/d(b1, b2).
template <bool LP>
__device__ __forceinline__ double eval(const Hyper& h, double b1, double b2, const double* __restrict__ y,
                                       const double* __restrict__ xs, int N, int Tn, bool has_rat,
                                       double (&gh)[6], double& g1, double& g2) {
  const double is2 = exp(-2.0 * h.ls);
  const double z = tanh(h.yo), D = 1.0 - z * z;
  const double it1 = 1.0 / h.t1, it2 = 1.0 / h.t2, iD = 1.0 / D;
  double sr = 0.0, srx = 0.0, ss = 0.0;
#pragma unroll
  for (int j = 0; j < TMAX; ++j) {
    if (j < Tn) {
      const double r = y[j] - fma(b2, xs[j], b1);
      sr += r; srx = fma(r, xs[j], srx); ss = fma(r, r, ss);
    }
  }
  const double a = (b1 - h.mu1) * it1, c = (b2 - h.mu2) * it2;
  const double da = (a - z * c) * iD, dc = (c - z * a) * iD;
  const double Q = (a * a - 2.0 * z * a * c + c * c) * iD;
  g1 = -da * it1 + sr * is2;
  g2 = -dc * it2 + srx * is2;
  const double w = has_rat ? 1.0 : 0.0;
  double s_mu1 = w * da * it1, s_mu2 = w * dc * it2;
  double s_t1 = w * (-it1 + a * da * it1), s_t2 = w * (-it2 + c * dc * it2);
  double s_z = w * (z * iD - 0.5 * ((-2.0 * a * c + 2.0 * z * Q) * iD));
  double s_ss = w * ss;
  s_mu1 = warp_sum(s_mu1); s_mu2 = warp_sum(s_mu2); s_t1 = warp_sum(s_t1); s_t2 = warp_sum(s_t2);
  s_z = warp_sum(s_z); s_ss = warp_sum(s_ss);
  const double sigma = exp(h.ls), zc = sigma * 0.4;            // sigma / 2.5
  gh[0] = s_mu1 - h.mu1 * 1e-4; gh[1] = s_mu2 - h.mu2 * 1e-4;
  gh[2] = s_t1 - 2.0 * h.t1 / (6.25 + h.t1 * h.t1); gh[3] = s_t2 - 2.0 * h.t2 / (6.25 + h.t2 * h.t2);
  gh[4] = (s_z - 4.0 * z * iD) * D;
  gh[5] = -(double)(N * Tn) + s_ss * is2 - 2.0 * zc * zc / (1.0 + zc * zc) + 1.0;
  double lp = 0.0;
  if (LP) {
    const double lD = log1p(-z * z);
    // per-rat part: multi_normal_lpdf(beta_i | betaMu, Sigma) + sum_j normal_lpdf(y_ij | ., sigma)
    double part = w * ((-kLog2Pi - log(h.t1) - log(h.t2) - 0.5 * lD - 0.5 * Q)
                       + (-(double)Tn * (kHalfLog2Pi + h.ls) - 0.5 * ss * is2));
    part = warp_sum(part);
    const double l100 = 4.6051701859880913680359829093687;     // log 100
    lp = (-kHalfLog2Pi - l100 - 0.5 * h.mu1 * h.mu1 * 1e-4) + (-kHalfLog2Pi - l100 - 0.5 * h.mu2 * h.mu2 * 1e-4);
    const double l25 = 0.91629073187415506518352721176801, l2 = 0.69314718055994530941723212145818;   // log 2.5, log 2
    lp += (-kLogPi - l25 - log1p(h.t1 * h.t1 * 0.16) + l2) + (-kLogPi - l25 - log1p(h.t2 * h.t2 * 0.16) + l2);
    lp += (lgamma(2.5) - (0.5 * kLogPi + lgamma(2.0))) + lD;  // lkj_corr_lpdf(Omega | 2), K = 2
    lp += lD;                                                  // Jacobian of tanh (corr_constrain.hpp:45-47)
    lp += (-kLogPi - l25 - log1p(zc * zc) + l2) + h.ls;       // sigma ~ cauchy(0, 2.5) T[0,] + lb Jacobian
    lp += part;
    if (h.t1 < 0.0 || h.t2 < 0.0) lp = neg_inf();             // T[0,] on the unconstrained tau of the generated program
  }
  return lp;
}

struct Ctx {
  int N, Tn, P, lane, wid;
  long long chain, cidx; bool active, has_rat;
  const double* xs;
};

__device__ __forceinline__ void setup(const StkArgs& a, double* smem, Ctx& c, double (&y)[TMAX]) {
  c.Tn = (int)a.flen[F_X];
  c.N = (int)(a.flen[F_Y] / max(c.Tn, 1));
  c.P = 2 * c.N + 6;
  c.lane = threadIdx.x & 31; c.wid = threadIdx.x >> 5;
  const double* X = (const double*)a.field[F_X];
  const double* Y = (const double*)a.field[F_Y];
  for (int i = threadIdx.x; i < c.Tn; i += blockDim.x) smem[i] = X[i];
  for (int i = threadIdx.x; i < c.N * c.Tn; i += blockDim.x) smem[TMAX + i] = Y[i];
  __syncthreads();
  c.xs = smem;
  c.has_rat = c.lane < c.N;
#pragma unroll
  for (int j = 0; j < TMAX; ++j) y[j] = (c.has_rat && j < c.Tn) ? smem[TMAX + c.lane * c.Tn + j] : 0.0;
  c.chain = (long long)blockIdx.x * (blockDim.x >> 5) + c.wid;
  c.active = c.chain < a.n_chains;
  c.cidx = c.active ? c.chain : a.n_chains - 1;
}

__device__ __forceinline__ void load_q(const StkArgs& a, const Ctx& c, Hyper& h, double& b1, double& b2) {
  const double* Q = (const double*)a.field[F_Q];
  const long long nc = a.n_chains;
  h.mu1 = Q[0 * nc + c.cidx]; h.mu2 = Q[1 * nc + c.cidx]; h.t1 = Q[2 * nc + c.cidx]; h.t2 = Q[3 * nc + c.cidx];
  h.yo = Q[4 * nc + c.cidx]; h.ls = Q[(long long)(c.P - 1) * nc + c.cidx];
  const int i = c.has_rat ? c.lane : 0;
  b1 = Q[(long long)(5 + i) * nc + c.cidx]; b2 = Q[(long long)(5 + c.N + i) * nc + c.cidx];
}

// one value of the Philox stream of the oracle's loop: p[j] = component (j & 1) of block j / 2
__device__ __forceinline__ double momentum(int j, double za, double zb, double za32, double zb32) {
  const int blk = j >> 1;
  const double va = __shfl_sync(0xffffffffu, za, blk & 31), vb = __shfl_sync(0xffffffffu, zb, blk & 31);
  const double v = (j & 1) ? vb : va;
  return blk < 32 ? v : ((j & 1) ? zb32 : za32);
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
// static HMC, n_steps transitions per launch
// ---------------------------------------------------------------------------------------------------------
extern "C" __global__ void __launch_bounds__(128) birats_hmc(const __grid_constant__ StkArgs a) {
  extern __shared__ __align__(16) double smem[];
  Ctx c; double y[TMAX];
  setup(a, smem, c, y);
  Hyper h; double b1, b2;
  load_q(a, c, h, b1, b2);
  const double eps = a.par[PAR_EPS];
  const int L = (int)a.par[PAR_L];
  const int j1 = 5 + (c.has_rat ? c.lane : 0), j2 = 5 + c.N + (c.has_rat ? c.lane : 0);
  unsigned long long n_acc = 0;
  double lp = 0.0;
  for (long long t = 0; t < a.n_steps; ++t) {
    Rng rng; rng.init(a.seed, (u64)(a.chain0 + c.cidx), a.step0 + (u64)t);
    // p ~ N(0, I): lane l draws block l; blocks 32.. (P > 64) by every lane (P = 66: one block)
    double za, zb, za32 = 0.0, zb32 = 0.0;
    { Rng r = rng; r.slot = (u32)c.lane; r.normal2(za, zb); }
    if (c.P > 64) { Rng r = rng; r.slot = 32u; r.normal2(za32, zb32); }
    double ph[6];
#pragma unroll
    for (int k = 0; k < 5; ++k) ph[k] = momentum(k, za, zb, za32, zb32);
    ph[5] = momentum(c.P - 1, za, zb, za32, zb32);
    double p1 = momentum(j1, za, zb, za32, zb32), p2 = momentum(j2, za, zb, za32, zb32);
    double gh[6], g1, g2;
    lp = eval<true>(h, b1, b2, y, c.xs, c.N, c.Tn, c.has_rat, gh, g1, g2);
    auto kinetic = [&]() {
      double ke = c.has_rat ? fma(p1, p1, 0.0) : 0.0;       // oracle order differs (sequential); 1e-16 relative
      ke = warp_sum(c.has_rat ? fma(p2, p2, ke) : 0.0);
#pragma unroll
      for (int k = 0; k < 6; ++k) ke = fma(ph[k], ph[k], ke);
      return ke;
    };
    const double H0 = -lp + 0.5 * kinetic();
    const Hyper h0 = h; const double b10 = b1, b20 = b2, lp0 = lp;
    for (int l = 0; l < L; ++l) {
      double* hv = &h.mu1;
#pragma unroll
      for (int k = 0; k < 6; ++k) { ph[k] = fma(0.5 * eps, gh[k], ph[k]); hv[k] = fma(eps, ph[k], hv[k]); }
      p1 = fma(0.5 * eps, g1, p1); b1 = fma(eps, p1, b1);
      p2 = fma(0.5 * eps, g2, p2); b2 = fma(eps, p2, b2);
      if (l + 1 < L) eval<false>(h, b1, b2, y, c.xs, c.N, c.Tn, c.has_rat, gh, g1, g2);
      else lp = eval<true>(h, b1, b2, y, c.xs, c.N, c.Tn, c.has_rat, gh, g1, g2);
#pragma unroll
      for (int k = 0; k < 6; ++k) ph[k] = fma(0.5 * eps, gh[k], ph[k]);
      p1 = fma(0.5 * eps, g1, p1); p2 = fma(0.5 * eps, g2, p2);
    }
    double H = -lp + 0.5 * kinetic();
    if (isnan(H)) H = pos_inf();
    const double accept_prob = exp(H0 - H);
    Rng ru = rng; ru.slot = (u32)((c.P + 1) / 2);
    const double u = ru.u01();
    if (accept_prob < 1.0 && u > accept_prob) { h = h0; b1 = b10; b2 = b20; lp = lp0; }
    else ++n_acc;
    long long slot;
    if (c.active && record_slot(a, t, &slot)) {
      if (c.lane == 0) {
        const double* hv = &h.mu1;
        for (int k = 0; k < 5; ++k) record_value(a, slot, k, c.chain, hv[k]);
        record_value(a, slot, c.P - 1, c.chain, h.ls);
      }
      if (c.has_rat) { record_value(a, slot, 5 + c.lane, c.chain, b1); record_value(a, slot, 5 + c.N + c.lane, c.chain, b2); }
    }
  }
  if (c.active) {
    double* Q = (double*)a.field[F_Q];
    const long long nc = a.n_chains;
    if (c.lane == 0) {
      Q[0 * nc + c.chain] = h.mu1; Q[1 * nc + c.chain] = h.mu2; Q[2 * nc + c.chain] = h.t1; Q[3 * nc + c.chain] = h.t2;
      Q[4 * nc + c.chain] = h.yo; Q[(long long)(c.P - 1) * nc + c.chain] = h.ls;
      ((double*)a.field[F_LP])[c.chain] = lp;
      if (a.accept) a.accept[c.chain] += n_acc;
    }
    if (c.has_rat) { Q[(long long)(5 + c.lane) * nc + c.chain] = b1; Q[(long long)(5 + c.N + c.lane) * nc + c.chain] = b2; }
  }
}

// ---------------------------------------------------------------------------------------------------------
// adaptive static HMC with a diagonal metric (Stan's adapt_diag_e_static_hmc per chain; stk_adapt.cuh,
// oracle/stk_oracle_adapt.cpp).  One warp = one chain, so every loop of the adaptation is warp-uniform; the metric
// and the Welford estimator are distributed like q (lane i: its rat's two coefficients; the six hyper-parameter
// entries replicated).  Philox blocks are consumed in the oracle's order: a momentum draw takes (P+1)/2 blocks from
// the running block counter, a uniform one.
// ---------------------------------------------------------------------------------------------------------
extern "C" __global__ void __launch_bounds__(128) birats_hmc_adapt(const __grid_constant__ StkArgs a) {
  extern __shared__ __align__(16) double smem[];
  Ctx c; double y[TMAX];
  setup(a, smem, c, y);
  Hyper h; double b1, b2;
  load_q(a, c, h, b1, b2);
  const long long nc = a.n_chains;
  const int r = c.has_rat ? c.lane : 0;
  const int j1 = 5 + r, j2 = 5 + c.N + r;
  // parameter index of hyper slot k
  auto hidx = [&](int k) { return k < 5 ? k : c.P - 1; };
  double* AD = (double*)a.field[F_AD];
  double* MINV = (double*)a.field[F_MINV];
  double* WM = (double*)a.field[F_WM];
  double* WM2 = (double*)a.field[F_WM2];
  AdaptCfg cfg; cfg.init(&a.par[PAR_ADAPT]);
  double ad[STK_AD_LEN], mh[6], m1, m2;
#pragma unroll
  for (int j = 0; j < STK_AD_LEN; ++j) ad[j] = AD[j * nc + c.cidx];
#pragma unroll
  for (int k = 0; k < 6; ++k) mh[k] = MINV[(long long)hidx(k) * nc + c.cidx];
  m1 = MINV[(long long)j1 * nc + c.cidx]; m2 = MINV[(long long)j2 * nc + c.cidx];
  const u32 blocks_p = (u32)((c.P + 1) / 2);
  unsigned long long n_acc = 0;
  double lp = 0.0, ph[6], p1, p2, gh[6], g1, g2;

  auto sample_p = [&](const Rng& rng, u32& base) {
    double za, zb, za32 = 0.0, zb32 = 0.0;
    { Rng q = rng; q.slot = base + (u32)c.lane; q.normal2(za, zb); }
    if (c.P > 64) { Rng q = rng; q.slot = base + 32u; q.normal2(za32, zb32); }
    base += blocks_p;
#pragma unroll
    for (int k = 0; k < 5; ++k) ph[k] = momentum(k, za, zb, za32, zb32) / sqrt(mh[k]);
    ph[5] = momentum(c.P - 1, za, zb, za32, zb32) / sqrt(mh[5]);
    p1 = momentum(j1, za, zb, za32, zb32) / sqrt(m1);
    p2 = momentum(j2, za, zb, za32, zb32) / sqrt(m2);
  };
  auto kinetic = [&]() {
    double ke = warp_sum(c.has_rat ? fma(m2 * p2, p2, m1 * p1 * p1) : 0.0);
#pragma unroll
    for (int k = 0; k < 6; ++k) ke = fma(mh[k] * ph[k], ph[k], ke);
    return 0.5 * ke;
  };
  // one leapfrog step of (h, b1, b2, p) with the gradient in (gh, g1, g2); returns lp at the new point
  auto evolve = [&](Hyper& hh, double& bb1, double& bb2, double eps) {
    double* hv = &hh.mu1;
#pragma unroll
    for (int k = 0; k < 6; ++k) { ph[k] = fma(0.5 * eps, gh[k], ph[k]); hv[k] = fma(eps * mh[k], ph[k], hv[k]); }
    p1 = fma(0.5 * eps, g1, p1); bb1 = fma(eps * m1, p1, bb1);
    p2 = fma(0.5 * eps, g2, p2); bb2 = fma(eps * m2, p2, bb2);
    const double l = eval<true>(hh, bb1, bb2, y, c.xs, c.N, c.Tn, c.has_rat, gh, g1, g2);
#pragma unroll
    for (int k = 0; k < 6; ++k) ph[k] = fma(0.5 * eps, gh[k], ph[k]);
    p1 = fma(0.5 * eps, g1, p1); p2 = fma(0.5 * eps, g2, p2);
    return l;
  };
  // base_hmc::init_stepsize at (h, b1, b2); lp and the gradient at that point are valid on entry and on exit
  auto init_stepsize = [&](const Rng& rng, u32& base) {
    double eps = ad[0];
    if (eps == 0.0 || eps > 1e7) return;
    const double lp0 = lp, g10 = g1, g20 = g2;
    double gh0[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) gh0[k] = gh[k];
    int dir = 0;
    for (;;) {
      Hyper ht = h; double bt1 = b1, bt2 = b2;
      sample_p(rng, base);
#pragma unroll
      for (int k = 0; k < 6; ++k) gh[k] = gh0[k];
      g1 = g10; g2 = g20;
      const double H0 = -lp0 + kinetic();
      const double lpt = evolve(ht, bt1, bt2, eps);
      double hh = -lpt + kinetic();
      if (isnan(hh)) hh = pos_inf();
      if (init_stepsize_update(eps, dir, H0 - hh)) break;
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) gh[k] = gh0[k];
    g1 = g10; g2 = g20;
    ad[0] = eps;
  };

  for (long long t = 0; t < a.n_steps; ++t) {
    const unsigned long long step = a.step0 + (u64)t;
    Rng rng; rng.init(a.seed, (u64)(a.chain0 + c.cidx), step);
    u32 base = 0;
    lp = eval<true>(h, b1, b2, y, c.xs, c.N, c.Tn, c.has_rat, gh, g1, g2);
    if (ad[9] == 0.0) {
#pragma unroll
      for (int k = 0; k < 6; ++k) mh[k] = 1.0;
      m1 = 1.0; m2 = 1.0;
      adapt_reset(ad, cfg);
      init_stepsize(rng, base);
    }
    if (ad[9] == 1.0 && (double)step >= cfg.num_warmup) { ad[0] = exp(ad[3]); ad[9] = 2.0; }
    const double eps = ad[0];
    const int L = adapt_leaps(cfg.int_time, eps);
    sample_p(rng, base);
    const double H0 = -lp + kinetic();
    const Hyper h0 = h; const double b10 = b1, b20 = b2, lp0 = lp;
    for (int l = 0; l < L; ++l) lp = evolve(h, b1, b2, eps);
    double H = -lp + kinetic();
    if (isnan(H)) H = pos_inf();
    double accept_prob = exp(H0 - H);
    Rng ru = rng; ru.slot = base; base += 1;
    const double u = ru.u01();
    if (accept_prob < 1.0 && u > accept_prob) { h = h0; b1 = b10; b2 = b20; lp = lp0; }
    else ++n_acc;
    if (accept_prob > 1.0) accept_prob = 1.0;
    if (ad[9] == 1.0) {
      adapt_learn_stepsize(ad, cfg, accept_prob);
      const bool in_w = adapt_in_window(ad, cfg), ends = adapt_window_ends(ad, cfg);
      bool update = false;
      if (in_w || ends) {
        const double* hv = &h.mu1;
        double wh[6], wh2[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) { wh[k] = WM[(long long)hidx(k) * nc + c.cidx]; wh2[k] = WM2[(long long)hidx(k) * nc + c.cidx]; }
        double w1 = WM[(long long)j1 * nc + c.cidx], w12 = WM2[(long long)j1 * nc + c.cidx];
        double w2 = WM[(long long)j2 * nc + c.cidx], w22 = WM2[(long long)j2 * nc + c.cidx];
        if (in_w) {
          ad[8] += 1.0;
#pragma unroll
          for (int k = 0; k < 6; ++k) welford_add(ad[8], hv[k], wh[k], wh2[k]);
          welford_add(ad[8], b1, w1, w12); welford_add(ad[8], b2, w2, w22);
        }
        if (ends) {
          adapt_next_window(ad, cfg);
#pragma unroll
          for (int k = 0; k < 6; ++k) { mh[k] = adapt_window_var(ad[8], wh2[k], mh[k]); wh[k] = 0.0; wh2[k] = 0.0; }
          m1 = adapt_window_var(ad[8], w12, m1); m2 = adapt_window_var(ad[8], w22, m2);
          w1 = w12 = w2 = w22 = 0.0;
          ad[8] = 0.0;
          update = true;
        }
        __syncwarp();   // all lanes have read the replicated hyper entries before lane 0 rewrites them
        if (c.active) {
          if (c.lane == 0) {
#pragma unroll
            for (int k = 0; k < 6; ++k) { WM[(long long)hidx(k) * nc + c.chain] = wh[k]; WM2[(long long)hidx(k) * nc + c.chain] = wh2[k]; }
          }
          if (c.has_rat) {
            WM[(long long)j1 * nc + c.chain] = w1; WM2[(long long)j1 * nc + c.chain] = w12;
            WM[(long long)j2 * nc + c.chain] = w2; WM2[(long long)j2 * nc + c.chain] = w22;
          }
        }
        __syncwarp();
      }
      ad[5] += 1.0;
      if (update) {
        lp = eval<true>(h, b1, b2, y, c.xs, c.N, c.Tn, c.has_rat, gh, g1, g2);
        init_stepsize(rng, base);
        ad[1] = log(10.0 * ad[0]); ad[2] = 0.0; ad[3] = 0.0; ad[4] = 0.0;
      }
    }
    long long slot;
    if (c.active && record_slot(a, t, &slot)) {
      if (c.lane == 0) {
        const double* hv = &h.mu1;
        for (int k = 0; k < 5; ++k) record_value(a, slot, k, c.chain, hv[k]);
        record_value(a, slot, c.P - 1, c.chain, h.ls);
      }
      if (c.has_rat) { record_value(a, slot, 5 + c.lane, c.chain, b1); record_value(a, slot, 5 + c.N + c.lane, c.chain, b2); }
    }
  }
  if (c.active) {
    double* Q = (double*)a.field[F_Q];
    if (c.lane == 0) {
      const double* hv = &h.mu1;
      for (int k = 0; k < 6; ++k) { Q[(long long)hidx(k) * nc + c.chain] = hv[k]; MINV[(long long)hidx(k) * nc + c.chain] = mh[k]; }
      for (int j = 0; j < STK_AD_LEN; ++j) AD[j * nc + c.chain] = ad[j];
      ((double*)a.field[F_LP])[c.chain] = lp;
      if (a.accept) a.accept[c.chain] += n_acc;
    }
    if (c.has_rat) {
      Q[(long long)j1 * nc + c.chain] = b1; Q[(long long)j2 * nc + c.chain] = b2;
      MINV[(long long)j1 * nc + c.chain] = m1; MINV[(long long)j2 * nc + c.chain] = m2;
    }
  }
}

template <bool GRAD>
__device__ __forceinline__ void logp_entry(const StkArgs& a) {
  extern __shared__ __align__(16) double smem[];
  Ctx c; double y[TMAX];
  setup(a, smem, c, y);
  Hyper h; double b1, b2;
  load_q(a, c, h, b1, b2);
  double gh[6], g1, g2;
  const double lp = eval<true>(h, b1, b2, y, c.xs, c.N, c.Tn, c.has_rat, gh, g1, g2);
  if (!c.active) return;
  const long long nc = a.n_chains;
  if (c.lane == 0) a.out_lp[c.chain] = lp;
  if (GRAD) {
    const bool ok = lp != neg_inf();      // Stan returns no gradient where the density is zero; report zeros
    if (c.lane == 0) {
      for (int k = 0; k < 5; ++k) a.out_grad[(long long)k * nc + c.chain] = ok ? gh[k] : 0.0;
      a.out_grad[(long long)(c.P - 1) * nc + c.chain] = ok ? gh[5] : 0.0;
    }
    if (c.has_rat) {
      a.out_grad[(long long)(5 + c.lane) * nc + c.chain] = ok ? g1 : 0.0;
      a.out_grad[(long long)(5 + c.N + c.lane) * nc + c.chain] = ok ? g2 : 0.0;
    }
  }
}
extern "C" __global__ void __launch_bounds__(128) birats_logp(const __grid_constant__ StkArgs a) { logp_entry<false>(a); }
extern "C" __global__ void __launch_bounds__(128) birats_logp_grad(const __grid_constant__ StkArgs a) { logp_entry<true>(a); }

// smem: x (TMAX doubles) + y (N * Tn doubles, 8 B per element of y)
extern "C" __device__ const char stk_desc[] =
    "stk-module 1\n"
    "name birats\n"
    "nparam 66\n"
    "field q state f64 66\n"
    "field lp state f64 1\n"
    "field x data f64 0\n"
    "field y data f64 0\n"
    "param eps 0.02\n"
    "field ad state f64 10\n"
    "field minv state f64 66\n"
    "field wm state f64 66\n"
    "field wm2 state f64 66\n"
    "param L 10\n"
    STK_ADAPT_PARAMS
    "record q\n"
    "kernel hmc birats_hmc 128 4 64 y 8\n"
    "kernel hmc_adapt birats_hmc_adapt 128 4 64 y 8\n"
    "entry logp birats_logp 128 4 64 y 8\n"
    "entry logp_grad birats_logp_grad 128 4 64 y 8\n";
