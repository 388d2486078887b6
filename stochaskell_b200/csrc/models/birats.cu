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
enum { F_Q = 0, F_LP = 1, F_X = 2, F_Y = 3, F_AD = 4, F_MINV = 5, F_WM = 6, F_WM2 = 7, F_NUTS = 8 };
enum { PAR_EPS = 0, PAR_L = 1, PAR_ADAPT = 2 /* stepsize .. window, stk_adapt.cuh */, PAR_MAXDEPTH = 12 };

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
  // lanes without a rat (N < 32) carry a phantom copy of rat 0 with no data: it must stay put, or a long adaptive launch
  // lets it run away under a metric estimated from its own wanderings until 0 * inf poisons the warp sums
  g1 = has_rat ? -da * it1 + sr * is2 : 0.0;
  g2 = has_rat ? -dc * it2 + srx * is2 : 0.0;
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
struct TreeNode { double rh[6], r1, r2, sh[6], s1, s2, qh[6], q1, q2, lsw, lp; };   // a pending left sibling (NUTS)
struct TreeEnd { double hv[6], b1, b2, ph[6], p1, p2, gh[6], g1, g2; };               // z_minus / z_plus

__device__ __forceinline__ double log_sum_exp2(double x, double y) {  // stan/math/prim/scal/fun/log_sum_exp.hpp
  if (x == neg_inf()) return y;
  if (x == pos_inf() && y == pos_inf()) return x;
  if (x > y) return x + log1p(exp(y - x));
  return y + log1p(exp(x - y));
}

// NUTS = false: adapt_diag_e_static_hmc;  NUTS = true: adapt_diag_e_nuts (base_nuts.hpp:71-244, the default of hmcStan —
// birats.ipynb cell 8 runs it), the tree built iteratively as in linreg_nuts; one warp = one chain, so the tree loops are
// warp-uniform and no state machine is needed.
template <bool NUTS>
__device__ __forceinline__ void birats_adaptive(const StkArgs& a) {
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
  int max_depth = (int)a.par[PAR_MAXDEPTH];
  max_depth = max_depth < 1 ? 1 : (max_depth > 10 ? 10 : max_depth);
  double st_depth = 0.0, st_leaps = 0.0, st_div = 0.0, st_acc = 0.0;
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
    if (!c.has_rat) { p1 = 0.0; p2 = 0.0; }
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
    double accept_prob;
    if constexpr (!NUTS) {
      const double eps = ad[0];
      const int L = adapt_leaps(cfg.int_time, eps);
      sample_p(rng, base);
      const double H0 = -lp + kinetic();
      const Hyper h0 = h; const double b10 = b1, b20 = b2, lp0 = lp;
      for (int l = 0; l < L; ++l) lp = evolve(h, b1, b2, eps);
      double H = -lp + kinetic();
      if (isnan(H)) H = pos_inf();
      accept_prob = exp(H0 - H);
      Rng ru = rng; ru.slot = base; base += 1;
      const double u = ru.u01();
      if (accept_prob < 1.0 && u > accept_prob) { h = h0; b1 = b10; b2 = b20; lp = lp0; }
      else ++n_acc;
    } else {
      const double eps = ad[0];
      sample_p(rng, base);
      auto save_end = [&](TreeEnd& e, const Hyper& zh, double zb1, double zb2) {
        const double* hv = &zh.mu1;
#pragma unroll
        for (int k = 0; k < 6; ++k) { e.hv[k] = hv[k]; e.ph[k] = ph[k]; e.gh[k] = gh[k]; }
        e.b1 = zb1; e.b2 = zb2; e.p1 = p1; e.p2 = p2; e.g1 = g1; e.g2 = g2;
      };
      auto dot = [&](const double* ah, double a1, double a2, const double* bh, double bb1, double bb2) {
        double d = warp_sum(c.has_rat ? fma(a2, bb2, a1 * bb1) : 0.0);
#pragma unroll
        for (int k = 0; k < 6; ++k) d = fma(ah[k], bh[k], d);
        return d;
      };
      TreeEnd ends2[2];
      save_end(ends2[0], h, b1, b2); ends2[1] = ends2[0];
      Hyper hs = h; double bs1 = b1, bs2 = b2, lps = lp;            // z_sample
      double rh[6], r1 = p1, r2 = p2;                                // rho
#pragma unroll
      for (int k = 0; k < 6; ++k) rh[k] = ph[k];
      double lsw = 0.0, sum_metro = 0.0;
      const double H0 = -lp + kinetic();
      int depth = 0, n_leap = 0;
      bool divergent = false;
      TreeNode st[10];
      while (depth < max_depth) {
        double ud; { Rng ru = rng; ru.slot = base; base += 1; ud = ru.u01(); }
        const int en = ud > 0.5 ? 1 : 0;
        const double e = en ? eps : -eps;
        Hyper zh; double zb1, zb2;
        {
          const TreeEnd& te = ends2[en];
          double* hv = &zh.mu1;
#pragma unroll
          for (int k = 0; k < 6; ++k) { hv[k] = te.hv[k]; ph[k] = te.ph[k]; gh[k] = te.gh[k]; }
          zb1 = te.b1; zb2 = te.b2; p1 = te.p1; p2 = te.p2; g1 = te.g1; g2 = te.g2;
        }
        TreeNode cur;
        bool valid = true;
        const int leaves = 1 << depth;
        for (int leaf = 0; leaf < leaves; ++leaf) {
          const double lpz = evolve(zh, zb1, zb2, e);
          ++n_leap;
          double hh = -lpz + kinetic();
          if (isnan(hh)) hh = pos_inf();
          if (hh - H0 > 1000.0) divergent = true;
          cur.lsw = H0 - hh; cur.lp = lpz;
          sum_metro += (H0 - hh > 0.0) ? 1.0 : exp(H0 - hh);
          const double* hv = &zh.mu1;
#pragma unroll
          for (int k = 0; k < 6; ++k) { cur.rh[k] = ph[k]; cur.sh[k] = mh[k] * ph[k]; cur.qh[k] = hv[k]; }
          cur.r1 = p1; cur.r2 = p2; cur.s1 = m1 * p1; cur.s2 = m2 * p2; cur.q1 = zb1; cur.q2 = zb2;
          valid = !divergent;
          int k = 0;
          while (valid && ((leaf >> k) & 1)) {
            const TreeNode& l = st[k];
            const double sub = log_sum_exp2(l.lsw, cur.lsw);
            bool take_right = cur.lsw > sub;
            if (!take_right) { Rng ru = rng; ru.slot = base; base += 1; take_right = ru.u01() < exp(cur.lsw - sub); }
            double psr[6];
#pragma unroll
            for (int j = 0; j < 6; ++j) {
              cur.rh[j] = l.rh[j] + cur.rh[j]; cur.sh[j] = l.sh[j]; psr[j] = mh[j] * ph[j];
              if (!take_right) cur.qh[j] = l.qh[j];
            }
            cur.r1 = l.r1 + cur.r1; cur.r2 = l.r2 + cur.r2; cur.s1 = l.s1; cur.s2 = l.s2;
            if (!take_right) { cur.q1 = l.q1; cur.q2 = l.q2; cur.lp = l.lp; }
            cur.lsw = sub;
            const double dr = dot(psr, m1 * p1, m2 * p2, cur.rh, cur.r1, cur.r2);
            const double dl = dot(cur.sh, cur.s1, cur.s2, cur.rh, cur.r1, cur.r2);
            valid = dr > 0.0 && dl > 0.0;
            ++k;
          }
          if (!valid) break;
          if (leaf + 1 < leaves) st[k] = cur;
        }
        if (!valid) break;
        save_end(ends2[en], zh, zb1, zb2);
        ++depth;
        bool take = cur.lsw > lsw;
        if (!take) { Rng ru = rng; ru.slot = base; base += 1; take = ru.u01() < exp(cur.lsw - lsw); }
        if (take) {
          double* hv = &hs.mu1;
#pragma unroll
          for (int k = 0; k < 6; ++k) hv[k] = cur.qh[k];
          bs1 = cur.q1; bs2 = cur.q2; lps = cur.lp;
        }
        lsw = log_sum_exp2(lsw, cur.lsw);
#pragma unroll
        for (int k = 0; k < 6; ++k) rh[k] += cur.rh[k];
        r1 += cur.r1; r2 += cur.r2;
        double sp[6], sm6[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) { sp[k] = mh[k] * ends2[1].ph[k]; sm6[k] = mh[k] * ends2[0].ph[k]; }
        const double dp = dot(sp, m1 * ends2[1].p1, m2 * ends2[1].p2, rh, r1, r2);
        const double dm = dot(sm6, m1 * ends2[0].p1, m2 * ends2[0].p2, rh, r1, r2);
        if (!(dp > 0.0 && dm > 0.0)) break;
      }
      accept_prob = sum_metro / (double)n_leap;
      h = hs; b1 = bs1; b2 = bs2; lp = lps;
      ++n_acc;
      st_depth = depth; st_leaps += n_leap; st_div += divergent ? 1.0 : 0.0; st_acc = accept_prob;
    }
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
      if (NUTS) {
        double* NS = (double*)a.field[F_NUTS];
        NS[0 * nc + c.chain] = st_depth; NS[1 * nc + c.chain] += st_leaps; NS[2 * nc + c.chain] += st_div; NS[3 * nc + c.chain] = st_acc;
      }
    }
    if (c.has_rat) {
      Q[(long long)j1 * nc + c.chain] = b1; Q[(long long)j2 * nc + c.chain] = b2;
      MINV[(long long)j1 * nc + c.chain] = m1; MINV[(long long)j2 * nc + c.chain] = m2;
    }
  }
}
extern "C" __global__ void __launch_bounds__(128) birats_hmc_adapt(const __grid_constant__ StkArgs a) { birats_adaptive<false>(a); }
extern "C" __global__ void __launch_bounds__(128) birats_nuts(const __grid_constant__ StkArgs a) { birats_adaptive<true>(a); }

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
    "field nuts state f64 4\n"
    "param L 10\n"
    STK_ADAPT_PARAMS
    "param max_depth 10\n"
    "record q\n"
    "kernel hmc birats_hmc 128 4 64 y 8\n"
    "kernel hmc_adapt birats_hmc_adapt 128 4 64 y 8\n"
    "kernel nuts birats_nuts 128 4 64 y 8\n"
    "entry logp birats_logp 128 4 64 y 8\n"
    "entry logp_grad birats_logp_grad 128 4 64 y 8\n";
