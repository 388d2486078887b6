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

using namespace stk;

namespace {

constexpr int TMAX = 8;                 // observations per rat held in registers
enum { F_Q = 0, F_LP = 1, F_X = 2, F_Y = 3 };
enum { PAR_EPS = 0, PAR_L = 1 };

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
    "param L 10\n"
    "record q\n"
    "kernel hmc birats_hmc 128 4 64 y 8\n"
    "entry logp birats_logp 128 4 64 y 8\n"
    "entry logp_grad birats_logp_grad 128 4 64 y 8\n";
