// stk_oracle_soccer.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// Restatement of the soccer reversible-jump kernel   soccer n `rjmcC` soccerJump   (rj.ipynb cells 1-10),
// node by node from the kernel IR that Language.CC embeds as the leading comment of the generated program
// (CC.hs:400-402; v_1_N names below are that IR's), with the statement templates of ccNode (CC.hs:123-274):
//   gamma_lpdf x a b        -> log pdf of gamma(shape a, scale 1/b)                  CC.hs:206-209
//   neg_binomial_lpdf k a b -> log pmf of negative_binomial(r = a, p = b/(b+1))      CC.hs:210-213
//   poisson_lpdf / normal_lpdf via the generic rule                                   CC.hs:226-228
//   foldl -> sequential loop (CC.hs:236-262);  `+s` -> left-to-right sum in operand order (CC.hs:92-107)
// in log space (Abstract.hs:68,113-115,121,152).
// PINNED against the reference's own generated C++ built into oracle/_ref/rj_soccer (oracle/build_ref.sh):
// tests/golden/soccer_ref.json holds (state, proposal draw) -> acceptance probability known answers printed
// by that binary with 17 significant digits (oracle/make_soccer_golden.py).
#include <cmath>
#include <cstdint>
#include "stk_oracle.h"

extern "C" {

// Model1 (tag 0): state (lam, y);  Model2 (tag 1): state (lam, kappa, y).
// `x` is the proposal's auxiliary draw x_cc_1_0 ~ normal 0 1.5 (used by the 0 -> 1 jump only).
// Returns the acceptance probability min(1, ratio) (v_1_19 / v_1_18) and the proposed kappa.
double orc_soccer_accept_prob(int tag, double lam, double kappa, const int* y, int ny, double x, double* kappa_prop) {
  const double lhalf = -0.6931471805599453;     // log(1/2): the model prior, printed so in the IR
  const double l015 = -4.199705077879927;       // log 0.015
  if (tag == 0) {
    const double v1 = orc_gamma_lpdf(lam, 25.0, 10.0);
    const double v2 = std::exp(x);
    const double v3 = 1.5e-2 * v2;               // kappa'
    const double v4 = orc_gamma_lpdf(v3, 1.0, 10.0);
    const double v5 = 1.0 / v3;
    const double v6 = v5 / lam;
    double v9 = 0.0;
    for (int i = 0; i < ny; ++i) v9 = v9 + orc_neg_binomial_lpdf(y[i], v5, v6);
    const double v10 = ((lhalf + v1) + v4) + v9;
    double v11 = 0.0;
    for (int i = 0; i < ny; ++i) v11 = v11 + orc_poisson_lpdf(y[i], lam);
    const double v12 = (lhalf + v1) + v11;
    const double v13 = v10 - v12;
    const double v14 = orc_normal_lpdf(x, 0.0, 1.5);
    const double v15 = -v14;
    const double v17 = ((x + v13) + v15) + l015;
    if (kappa_prop) *kappa_prop = v3;
    return std::fmin(1.0, std::exp(v17));
  }
  const double v0 = orc_gamma_lpdf(lam, 25.0, 10.0);
  double v3 = 0.0;
  for (int i = 0; i < ny; ++i) v3 = v3 + orc_poisson_lpdf(y[i], lam);
  const double v4 = (lhalf + v0) + v3;
  const double v5 = orc_gamma_lpdf(kappa, 1.0, 10.0);
  const double v6 = 1.0 / kappa;
  const double v7 = v6 / lam;
  double v8 = 0.0;
  for (int i = 0; i < ny; ++i) v8 = v8 + orc_neg_binomial_lpdf(y[i], v6, v7);
  const double v9 = ((lhalf + v0) + v5) + v8;
  const double v10 = v4 - v9;
  const double v11 = std::log(kappa);
  const double v12 = v11 - l015;
  const double v13 = orc_normal_lpdf(v12, 0.0, 1.5);
  const double v14 = -v11;
  const double v16 = (v10 + v13) + v14;
  if (kappa_prop) *kappa_prop = kappa;
  return std::fmin(1.0, std::exp(v16));
}

// one transition with the Philox layout of stkc-generated kernels (csrc/stkc.cpp): the PNode x_<ns>_<L>_<P> draws with
// counter word c2 ^ (L << 24 | P << 12), blocks counted from 0.  Model1 alternative: x_cc_1_0 ~ normal (L 1, P 0),
// x_cc_1_1 ~ bernoulli (L 1, P 1); Model2 alternative: x_cc_1_0 ~ bernoulli (L 1, P 0).  Returns the new tag; *kappa
// updated on a 0 -> 1 jump.
int orc_soccer_rj_step(int tag, double lam, double* kappa, const int* y, int ny,
                       uint64_t seed, uint64_t chain, uint64_t step, double* p_out) {
  orc_rng g0; orc_rng_init(&g0, seed, chain, step);
  const uint32_t t10 = (1u << 24) | (0u << 12), t11 = (1u << 24) | (1u << 12);
  orc_rng g = g0;
  double x = 0.0;
  if (tag == 0) { g.c2 ^= t10; g.slot = 0; x = 0.0 + 1.5 * orc_normal(&g); }
  double kp = *kappa;
  const double p = orc_soccer_accept_prob(tag, lam, *kappa, y, ny, x, &kp);
  g = g0; g.c2 ^= tag == 0 ? t11 : t10; g.slot = 0;
  const int acc = orc_draw_bernoulli(&g, p);
  if (p_out) *p_out = p;
  if (!acc) return tag;
  if (tag == 0) { *kappa = kp; return 1; }
  return 0;
}

}  // extern "C"
