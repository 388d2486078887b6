// stk_hmc_generic.cuh — the adaptive samplers behind `hmcStan` for GENERATED modules (csrc/stkc_grad.inc): Stan's
// adapt_diag_e_static_hmc and adapt_diag_e_nuts (the DEFAULT of hmcStan, Stan.hs:428-440) written once over an evaluation
// functor, so that a model lowered from IR gets the same samplers as the hand-tuned builtins (models/linreg.cu:246-720 is the
// model-specific form of the same code; oracle/stk_oracle_adapt.cpp the step-by-step restatement both are tested against).
//
//   ev(q, g) -> lp      log-density and gradient at q (unconstrained scale).  In a module with a sum loop the K observation-slice
//                       threads of a chain call it together (it contains CTA barriers), so every loop whose trip count differs
//                       between chains runs to the CTA's maximum: any(b) = __syncthreads_or(b) there, b itself otherwise.
//   State per chain: q, lp, ad[STK_AD_LEN], minv[P], wm[P], wm2[P], nuts[4] (fields of the module), replicated in the slice
//   threads of a chain, which take every decision redundantly on identical numbers.
// Philox use per transition is that of the builtins (momenta: one block per pair; then the uniforms of the sampler in the order
// the reference draws them), so a generated and a hand-tuned module of one model follow each other from the same seed.
#ifndef STK_HMC_GENERIC_CUH
#define STK_HMC_GENERIC_CUH
#include "stk_adapt.cuh"

namespace stk {

// loops over the P coordinates are unrolled (state in registers) for small models only: a 66-parameter model would be compiled for minutes
#define STK_GEN_UNROLL _Pragma("unroll (P <= 16 ? P : 1)")

struct GenFields { double *Q, *LP, *AD, *MINV, *WM, *WM2, *NS; };

STK_DEV double gen_log_sum_exp2(double x, double y) {  // stan/math/prim/scal/fun/log_sum_exp.hpp
  if (x == neg_inf()) return y;
  if (x == pos_inf() && y == pos_inf()) return x;
  if (x > y) return x + log1p(exp(y - x));
  return y + log1p(exp(x - y));
}

// ---- adapt_diag_e_static_hmc (adapt_diag_e_static_hmc.hpp:29-51, base_static_hmc.hpp:42-72) ------------------------------------
template <int P, class Eval, class Any>
STK_DEV void gen_hmc_adapt(const StkArgs& a, const GenFields& f, const double* par_adapt, long long chain, long long cidx, bool writer,
                           Eval&& ev, Any&& any) {
  const long long nc = a.n_chains;
  AdaptCfg cfg; cfg.init(par_adapt);
  double q[P], g[P], p[P], minv[P], ad[STK_AD_LEN];
  STK_GEN_UNROLL for (int j = 0; j < P; ++j) { q[j] = f.Q[j * nc + cidx]; minv[j] = f.MINV[j * nc + cidx]; }
  for (int j = 0; j < STK_AD_LEN; ++j) ad[j] = f.AD[j * nc + cidx];
  unsigned long long n_acc = 0;
  double lp = 0.0;
  auto sample_p = [&](Rng& rng) {
    STK_GEN_UNROLL for (int j = 0; j < P; j += 2) {
      double z0, z1; rng.normal2(z0, z1);
      p[j] = z0 / sqrt(minv[j]); if (j + 1 < P) p[j + 1] = z1 / sqrt(minv[j + 1]);
    }
  };
  auto kinetic = [&]() {
    double ke = 0.0;
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) ke = fma(minv[j] * p[j], p[j], ke);
    return 0.5 * ke;
  };
  // base_hmc::init_stepsize (base_hmc.hpp:73-126); lp / g hold the density and gradient at q on entry and on exit
  auto init_stepsize = [&](Rng& rng, bool want) {
    double eps = ad[0];
    int dir = 0;
    bool done = !want || eps == 0.0 || eps > 1e7;
    const double lp0 = lp;
    double g0[P], qt[P], gt[P];
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) g0[j] = g[j];
    while (any(!done)) {
      if (!done) sample_p(rng);
      const double H0 = -lp0 + kinetic();
      STK_GEN_UNROLL for (int j = 0; j < P; ++j) { p[j] = fma(0.5 * eps, g0[j], p[j]); qt[j] = fma(eps * minv[j], p[j], q[j]); }
      const double lpt = ev(qt, gt);
      STK_GEN_UNROLL for (int j = 0; j < P; ++j) p[j] = fma(0.5 * eps, gt[j], p[j]);
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
      STK_GEN_UNROLL for (int j = 0; j < P; ++j) minv[j] = 1.0;
      adapt_reset(ad, cfg);
    }
    lp = ev(q, g);
    if (any(fresh)) init_stepsize(rng, fresh);
    if (ad[9] == 1.0 && (double)step >= cfg.num_warmup) { ad[0] = exp(ad[3]); ad[9] = 2.0; }  // complete_adaptation
    const double eps = ad[0];
    const int L = adapt_leaps(cfg.int_time, eps);      // follows the step size in force (as the builtins; DESIGN.md 3, divergences)
    sample_p(rng);
    const double H0 = -lp + kinetic();
    double q0[P]; const double lp0 = lp;
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) q0[j] = q[j];
    for (int l = 0; any(l < L); ++l) {
      const bool on = l < L;
      if (on) { STK_GEN_UNROLL for (int j = 0; j < P; ++j) { p[j] = fma(0.5 * eps, g[j], p[j]); q[j] = fma(eps * minv[j], p[j], q[j]); } }
      lp = ev(q, g);                                   // a chain that is done evaluates at its final point again: same lp, same g
      if (on) { STK_GEN_UNROLL for (int j = 0; j < P; ++j) p[j] = fma(0.5 * eps, g[j], p[j]); }
    }
    double H = -lp + kinetic();
    if (isnan(H)) H = pos_inf();
    double accept_prob = exp(H0 - H);
    const double u = rng.u01();
    if (accept_prob < 1.0 && u > accept_prob) {
      STK_GEN_UNROLL for (int j = 0; j < P; ++j) q[j] = q0[j];
      lp = lp0;
    } else ++n_acc;
    if (accept_prob > 1.0) accept_prob = 1.0;
    bool update = false, store = false;
    double wm[P], wm2[P];
    if (ad[9] == 1.0) {
      adapt_learn_stepsize(ad, cfg, accept_prob);
      const bool in_w = adapt_in_window(ad, cfg), ends = adapt_window_ends(ad, cfg);
      if (in_w || ends) {
        // the Welford estimator lives in the module's wm / wm2 fields; the replicas of a chain compute identical values
        STK_GEN_UNROLL for (int j = 0; j < P; ++j) { wm[j] = f.WM[j * nc + cidx]; wm2[j] = f.WM2[j * nc + cidx]; }
        if (in_w) {
        ad[8] += 1.0;
        STK_GEN_UNROLL for (int j = 0; j < P; ++j) welford_add(ad[8], q[j], wm[j], wm2[j]);
      }
        if (ends) {
          adapt_next_window(ad, cfg);
          STK_GEN_UNROLL for (int j = 0; j < P; ++j) { minv[j] = adapt_window_var(ad[8], wm2[j], minv[j]); wm[j] = 0.0; wm2[j] = 0.0; }
          ad[8] = 0.0;
          update = true;
        }
        store = true;
      }
      ad[5] += 1.0;
    }
    any(false);                                        // every replica has read the estimator before the writer stores it
    if (store && writer) { STK_GEN_UNROLL for (int j = 0; j < P; ++j) { f.WM[j * nc + chain] = wm[j]; f.WM2[j * nc + chain] = wm2[j]; } }
    if (any(update)) {
      lp = ev(q, g);
      init_stepsize(rng, update);
      if (update) { ad[1] = log(10.0 * ad[0]); ad[2] = 0.0; ad[3] = 0.0; ad[4] = 0.0; }
    }
    long long slot;
    if (writer && record_slot(a, t, &slot)) { STK_GEN_UNROLL for (int j = 0; j < P; ++j) record_value(a, slot, j, chain, q[j]); }
  }
  if (writer) {
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) { f.Q[j * nc + chain] = q[j]; f.MINV[j * nc + chain] = minv[j]; }
    for (int j = 0; j < STK_AD_LEN; ++j) f.AD[j * nc + chain] = ad[j];
    f.LP[chain] = lp;
    if (a.accept) a.accept[chain] += n_acc;
  }
}

// ---- adapt_diag_e_nuts (base_nuts.hpp:71-244, 2.20 form: multinomial sampling, generalised U-turn on rho; adapt_diag_e_nuts.hpp:27-48)
// The recursive build_tree is iterative (leaf i of a subtree is followed by one merge per trailing 1 bit of i against a stack of
// pending left siblings: the same pairings and uniforms in the same post-order) and every chain runs its own state machine whose
// tick is one gradient evaluation; a chain that finishes a transition starts its next one at once.
constexpr int GEN_NUTS_MAXD = 10;
enum { GPH_START = 0, GPH_TRIAL = 1, GPH_LEAF = 2, GPH_UPD = 3, GPH_DONE = 4 };

template <int P, class Eval, class Any>
STK_DEV void gen_nuts(const StkArgs& a, const GenFields& f, const double* par_adapt, double max_depth_par, long long chain, long long cidx,
                      bool writer, Eval&& ev, Any&& any) {
  const long long nc = a.n_chains;
  AdaptCfg cfg; cfg.init(par_adapt);
  int max_depth = (int)max_depth_par;
  max_depth = max_depth < 1 ? 1 : (max_depth > GEN_NUTS_MAXD ? GEN_NUTS_MAXD : max_depth);
  double q[P], minv[P], wm[P], wm2[P], ad[STK_AD_LEN];
  double qe[P], ge[P];                          // evaluation point of the next tick / its gradient
  double g0[P], lp0 = 0.0;                      // density and gradient at q
  double zq[P], zp[P], zg[P];                   // the end of the trajectory that is being extended
  double eq[2][P], ep[2][P], eg[2][P];          // z_minus (0) and z_plus (1)
  double qs[P], lps = 0.0, rho[P];              // z_sample, rho of the whole tree
  double st_rho[GEN_NUTS_MAXD][P], st_psl[GEN_NUTS_MAXD][P], st_q[GEN_NUTS_MAXD][P], st_lsw[GEN_NUTS_MAXD], st_lp[GEN_NUTS_MAXD];
  double lp = 0.0, lsw = 0.0, H0 = 0.0, sum_metro = 0.0, eps = 0.0, H0t = 0.0, teps = 0.0;
  double stat_depth = 0.0, stat_leaps = 0.0, stat_div = 0.0, stat_acc = 0.0;
  int phase = a.n_steps > 0 ? GPH_START : GPH_DONE, depth = 0, leaf = 0, sgn = 1, n_leap = 0, tdir = 0;
  bool divergent = false, from_update = false;
  long long t = 0;
  unsigned long long n_acc = 0;
  Rng rng; rng.init(a.seed, (u64)(a.chain0 + cidx), a.step0);
  STK_GEN_UNROLL for (int j = 0; j < P; ++j) {
    q[j] = f.Q[j * nc + cidx]; minv[j] = f.MINV[j * nc + cidx]; wm[j] = f.WM[j * nc + cidx]; wm2[j] = f.WM2[j * nc + cidx];
    qe[j] = q[j];
  }
  for (int j = 0; j < STK_AD_LEN; ++j) ad[j] = f.AD[j * nc + cidx];

  auto kinetic = [&](const double* p) {
    double ke = 0.0;
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) ke = fma(minv[j] * p[j], p[j], ke);
    return 0.5 * ke;
  };
  auto sample_p = [&](double* p) {
    STK_GEN_UNROLL for (int j = 0; j < P; j += 2) {
      double z0, z1; rng.normal2(z0, z1);
      p[j] = z0 / sqrt(minv[j]); if (j + 1 < P) p[j + 1] = z1 / sqrt(minv[j + 1]);
    }
  };
  auto half_step = [&](const double* q_from, const double* g_from, double e) {
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) { zp[j] = fma(0.5 * e, g_from[j], zp[j]); qe[j] = fma(e * minv[j], zp[j], q_from[j]); }
  };
  auto prepare_leaf = [&]() { half_step(zq, zg, sgn * eps); phase = GPH_LEAF; };
  auto start_subtree = [&]() {
    sgn = rng.u01() > 0.5 ? 1 : -1;
    const int e = sgn > 0 ? 1 : 0;
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) { zq[j] = eq[e][j]; zp[j] = ep[e][j]; zg[j] = eg[e][j]; }
    leaf = 0;
    prepare_leaf();
  };
  auto prepare_trial = [&]() { sample_p(zp); H0t = -lp0 + kinetic(zp); half_step(q, g0, teps); phase = GPH_TRIAL; };
  auto begin_transition = [&]() {
    const unsigned long long step = a.step0 + (u64)t;
    if (ad[9] == 1.0 && (double)step >= cfg.num_warmup) { ad[0] = exp(ad[3]); ad[9] = 2.0; }  // complete_adaptation
    eps = ad[0];
    sample_p(zp);
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) {
      eq[0][j] = eq[1][j] = q[j]; ep[0][j] = ep[1][j] = zp[j]; eg[0][j] = eg[1][j] = g0[j];
      qs[j] = q[j]; rho[j] = zp[j];
    }
    lps = lp0; lsw = 0.0; H0 = -lp0 + kinetic(zp);
    depth = 0; n_leap = 0; sum_metro = 0.0; divergent = false;
    start_subtree();
  };
  auto begin_init_stepsize = [&](bool upd) {
    from_update = upd; teps = ad[0]; tdir = 0;
    if (teps == 0.0 || teps > 1e7) return false;
    prepare_trial();
    return true;
  };
  auto next_transition = [&]() {
    ++t;
    if (t >= a.n_steps) { phase = GPH_DONE; return; }
    rng.init(a.seed, (u64)(a.chain0 + cidx), a.step0 + (u64)t);
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) qe[j] = q[j];
    phase = GPH_START;
  };
  auto end_transition = [&]() {
    const double accept_prob = sum_metro / (double)n_leap;
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) q[j] = qs[j];
    lp = lps;
    ++n_acc;
    stat_depth = depth; stat_leaps += n_leap; stat_div += divergent ? 1.0 : 0.0; stat_acc = accept_prob;
    long long slot;
    if (writer && record_slot(a, t, &slot)) { STK_GEN_UNROLL for (int j = 0; j < P; ++j) record_value(a, slot, j, chain, q[j]); }
    bool update = false;
    if (ad[9] == 1.0) {
      adapt_learn_stepsize(ad, cfg, accept_prob > 1.0 ? 1.0 : accept_prob);
      const bool in_w = adapt_in_window(ad, cfg), ends = adapt_window_ends(ad, cfg);
      if (in_w) {
        ad[8] += 1.0;
        STK_GEN_UNROLL for (int j = 0; j < P; ++j) welford_add(ad[8], q[j], wm[j], wm2[j]);
      }
      if (ends) {
        adapt_next_window(ad, cfg);
        STK_GEN_UNROLL for (int j = 0; j < P; ++j) { minv[j] = adapt_window_var(ad[8], wm2[j], minv[j]); wm[j] = 0.0; wm2[j] = 0.0; }
        ad[8] = 0.0;
        update = true;
      }
      ad[5] += 1.0;
    }
    if (update) {
      STK_GEN_UNROLL for (int j = 0; j < P; ++j) qe[j] = q[j];
      phase = GPH_UPD;
    }
    else next_transition();
  };

  for (;;) {
    const bool busy = phase != GPH_DONE;
    if (!any(busy)) break;
    const double lpe = ev(qe, ge);
    if (!busy) continue;
    if (phase == GPH_START || phase == GPH_UPD) {
      lp0 = lpe; lp = lpe;
      STK_GEN_UNROLL for (int j = 0; j < P; ++j) g0[j] = ge[j];
      if (phase == GPH_START) {
        bool trials = false;
        if (ad[9] == 0.0) {
          STK_GEN_UNROLL for (int j = 0; j < P; ++j) { minv[j] = 1.0; wm[j] = 0.0; wm2[j] = 0.0; }
          adapt_reset(ad, cfg);
          trials = begin_init_stepsize(false);
        }
        if (!trials) begin_transition();
      } else if (!begin_init_stepsize(true)) {
        ad[1] = log(10.0 * ad[0]); ad[2] = 0.0; ad[3] = 0.0; ad[4] = 0.0;
        next_transition();
      }
    } else if (phase == GPH_TRIAL) {
      STK_GEN_UNROLL for (int j = 0; j < P; ++j) zp[j] = fma(0.5 * teps, ge[j], zp[j]);
      double h = -lpe + kinetic(zp);
      if (isnan(h)) h = pos_inf();
      if (!init_stepsize_update(teps, tdir, H0t - h)) prepare_trial();
      else {
        ad[0] = teps;
        if (from_update) { ad[1] = log(10.0 * ad[0]); ad[2] = 0.0; ad[3] = 0.0; ad[4] = 0.0; next_transition(); }
        else begin_transition();
      }
    } else {  // GPH_LEAF: build_tree(depth == 0), then the merges that complete with this leaf
      const double e = sgn * eps;
      STK_GEN_UNROLL for (int j = 0; j < P; ++j) { zq[j] = qe[j]; zg[j] = ge[j]; zp[j] = fma(0.5 * e, ge[j], zp[j]); }
      ++n_leap;
      double h = -lpe + kinetic(zp);
      if (isnan(h)) h = pos_inf();
      if (h - H0 > 1000.0) divergent = true;
      double c_lsw = H0 - h, c_lp = lpe;
      sum_metro += (H0 - h > 0.0) ? 1.0 : exp(H0 - h);
      double c_rho[P], c_psl[P], c_q[P];
      STK_GEN_UNROLL for (int j = 0; j < P; ++j) { c_rho[j] = zp[j]; c_psl[j] = minv[j] * zp[j]; c_q[j] = zq[j]; }
      bool valid = !divergent;
      int k = 0;
      while (valid && ((leaf >> k) & 1)) {
        const double l_lsw = st_lsw[k];
        const double sub = gen_log_sum_exp2(l_lsw, c_lsw);
        bool take_right = c_lsw > sub;
        if (!take_right) take_right = rng.u01() < exp(c_lsw - sub);
        double dl = 0.0, dr = 0.0;
        STK_GEN_UNROLL for (int j = 0; j < P; ++j) {
          const double r = st_rho[k][j] + c_rho[j];
          c_rho[j] = r;
          c_psl[j] = st_psl[k][j];
          if (!take_right) c_q[j] = st_q[k][j];
          dr = fma(minv[j] * zp[j], r, dr); dl = fma(c_psl[j], r, dl);
        }
        if (!take_right) c_lp = st_lp[k];
        c_lsw = sub;
        valid = dr > 0.0 && dl > 0.0;
        ++k;
      }
      if (!valid) end_transition();
      else if (leaf + 1 == (1 << depth)) {
        const int en = sgn > 0 ? 1 : 0;
        STK_GEN_UNROLL for (int j = 0; j < P; ++j) { eq[en][j] = zq[j]; ep[en][j] = zp[j]; eg[en][j] = zg[j]; }
        ++depth;
        bool take = c_lsw > lsw;
        if (!take) take = rng.u01() < exp(c_lsw - lsw);
        lsw = gen_log_sum_exp2(lsw, c_lsw);
        double dp = 0.0, dm = 0.0;
        STK_GEN_UNROLL for (int j = 0; j < P; ++j) {
          if (take) qs[j] = c_q[j];
          rho[j] += c_rho[j];
          dp = fma(minv[j] * ep[1][j], rho[j], dp); dm = fma(minv[j] * ep[0][j], rho[j], dm);
        }
        if (take) lps = c_lp;
        if (!(dp > 0.0 && dm > 0.0) || depth >= max_depth) end_transition();
        else start_subtree();
      } else {
        STK_GEN_UNROLL for (int j = 0; j < P; ++j) { st_rho[k][j] = c_rho[j]; st_psl[k][j] = c_psl[j]; st_q[k][j] = c_q[j]; }
        st_lsw[k] = c_lsw; st_lp[k] = c_lp;
        ++leaf;
        prepare_leaf();
      }
    }
  }
  if (writer) {
    STK_GEN_UNROLL for (int j = 0; j < P; ++j) {
      f.Q[j * nc + chain] = q[j]; f.MINV[j * nc + chain] = minv[j]; f.WM[j * nc + chain] = wm[j]; f.WM2[j * nc + chain] = wm2[j];
    }
    for (int j = 0; j < STK_AD_LEN; ++j) f.AD[j * nc + chain] = ad[j];
    f.NS[0 * nc + chain] = stat_depth; f.NS[1 * nc + chain] += stat_leaps; f.NS[2 * nc + chain] += stat_div; f.NS[3 * nc + chain] = stat_acc;
    f.LP[chain] = lp;
    if (a.accept) a.accept[chain] += n_acc;
  }
}

// ---- adapt_diag_e_nuts with the chain's state DISTRIBUTED over the G lanes of a chain (lanes modules of the gradient lowering) --------
// The same sampler as gen_nuts — same phases, same Philox blocks, same uniforms in the same order — for a CTA of one warp = CPB chains
// x G lanes.  The 21 P-vectors of a chain (position, metric, Welford estimator, evaluation point and gradient, trajectory ends, sample,
// rho and the candidate of the subtree being built) live in SHARED memory [array][j][chain in CTA]; coordinate pair (2b, 2b + 1)
// belongs to lane b mod G (the ownership of the lane-parallel evaluation), which does all element-wise work on it; dot products are
// partial sums combined by an xor butterfly over the chain's lanes (bitwise identical on them, so every lane takes the same
// decision); scalars and the adaptation state are replicated in registers; the stack of pending left siblings is distributed too
// (each lane keeps its own coordinates, in local memory: it is touched once or twice per leapfrog).  Chains of one warp sit in different
// phases, so everything inside a phase uses the chain's own lane mask; the evaluation is the one warp-wide step of a tick.
//   ev() -> lp   evaluates at the array QE (the evaluation reads it through its position mirror) and leaves each lane's gradient
//                entries in GE; called by all 32 lanes together.
enum { GA_Q = 0, GA_MINV, GA_WM, GA_WM2, GA_QE, GA_GE, GA_G0, GA_ZQ, GA_ZP, GA_ZG, GA_EQ0, GA_EQ1, GA_EP0, GA_EP1, GA_EG0, GA_EG1, GA_QS, GA_RHO,
       GA_CRHO, GA_CPSL, GA_CQ, GA_COUNT };

template <int P, int G, int CPB, class Eval>
STK_DEV void gen_nuts_lanes(const StkArgs& a, const GenFields& f, const double* par_adapt, double max_depth_par, long long chain, long long cidx,
                            bool active, int slice, int c, double* sm, Eval&& ev) {
  constexpr int NPAIR = (P + 1) / 2, NO = 2 * ((NPAIR + G - 1) / G);
  const long long nc = a.n_chains;
  const unsigned gmask = (G == 32 ? 0xffffffffu : ((1u << G) - 1u)) << ((threadIdx.x / G) * G);
  const bool writer = slice == 0 && active;
  AdaptCfg cfg; cfg.init(par_adapt);
  int max_depth = (int)max_depth_par;
  max_depth = max_depth < 1 ? 1 : (max_depth > GEN_NUTS_MAXD ? GEN_NUTS_MAXD : max_depth);
#define GAT(arr, j) sm[((arr) * P + (j)) * CPB + c]
#define GOWN(k) (2 * (slice + ((k) >> 1) * G) + ((k) & 1))
#define GEACH(k, j) for (int k = 0; k < NO; ++k) if (const int j = GOWN(k); j < P)
  auto gsum = [&](double v) { return group_sum_masked<G>(v, gmask); };
  double ad[STK_AD_LEN];
  double st_rho[GEN_NUTS_MAXD][NO], st_psl[GEN_NUTS_MAXD][NO], st_q[GEN_NUTS_MAXD][NO], st_lsw[GEN_NUTS_MAXD], st_lp[GEN_NUTS_MAXD];
  double lp0 = 0.0, lps = 0.0, lp = 0.0, lsw = 0.0, H0 = 0.0, sum_metro = 0.0, eps = 0.0, H0t = 0.0, teps = 0.0;
  double stat_depth = 0.0, stat_leaps = 0.0, stat_div = 0.0, stat_acc = 0.0;
  int phase = a.n_steps > 0 ? GPH_START : GPH_DONE, depth = 0, leaf = 0, sgn = 1, n_leap = 0, tdir = 0;
  bool divergent = false, from_update = false;
  long long t = 0;
  unsigned long long n_acc = 0;
  Rng rng; rng.init(a.seed, (u64)(a.chain0 + cidx), a.step0);
  GEACH(k, j) {
    GAT(GA_Q, j) = f.Q[j * nc + cidx]; GAT(GA_MINV, j) = f.MINV[j * nc + cidx]; GAT(GA_WM, j) = f.WM[j * nc + cidx]; GAT(GA_WM2, j) = f.WM2[j * nc + cidx];
    GAT(GA_QE, j) = GAT(GA_Q, j);
  }
  for (int j = 0; j < STK_AD_LEN; ++j) ad[j] = f.AD[j * nc + cidx];

  auto kinetic = [&](int parr) { double ke = 0.0; GEACH(k, j) ke = fma(GAT(GA_MINV, j) * GAT(parr, j), GAT(parr, j), ke); return 0.5 * gsum(ke); };
  auto sample_p = [&](int parr) {                       // pair b of the momenta is Philox block (current slot + b), drawn by its owner
    const u32 s0 = rng.slot;
    for (int kk = 0; kk < NO / 2; ++kk) {
      const int b = slice + kk * G, j = 2 * b;
      if (j < P) {
        rng.slot = s0 + (u32)b;
        double z0, z1; rng.normal2(z0, z1);
        GAT(parr, j) = z0 / sqrt(GAT(GA_MINV, j)); if (j + 1 < P) GAT(parr, j + 1) = z1 / sqrt(GAT(GA_MINV, j + 1));
      }
    }
    rng.slot = s0 + (u32)NPAIR;
  };
  auto half_step = [&](int qfrom, int gfrom, double e) {
    GEACH(k, j) { GAT(GA_ZP, j) = fma(0.5 * e, GAT(gfrom, j), GAT(GA_ZP, j)); GAT(GA_QE, j) = fma(e * GAT(GA_MINV, j), GAT(GA_ZP, j), GAT(qfrom, j)); }
  };
  auto prepare_leaf = [&]() { half_step(GA_ZQ, GA_ZG, sgn * eps); phase = GPH_LEAF; };
  auto start_subtree = [&]() {
    sgn = rng.u01() > 0.5 ? 1 : -1;
    const int eq = sgn > 0 ? GA_EQ1 : GA_EQ0, ep = sgn > 0 ? GA_EP1 : GA_EP0, eg = sgn > 0 ? GA_EG1 : GA_EG0;
    GEACH(k, j) { GAT(GA_ZQ, j) = GAT(eq, j); GAT(GA_ZP, j) = GAT(ep, j); GAT(GA_ZG, j) = GAT(eg, j); }
    leaf = 0;
    prepare_leaf();
  };
  auto prepare_trial = [&]() { sample_p(GA_ZP); H0t = -lp0 + kinetic(GA_ZP); half_step(GA_Q, GA_G0, teps); phase = GPH_TRIAL; };
  auto begin_transition = [&]() {
    const unsigned long long step = a.step0 + (u64)t;
    if (ad[9] == 1.0 && (double)step >= cfg.num_warmup) { ad[0] = exp(ad[3]); ad[9] = 2.0; }  // complete_adaptation
    eps = ad[0];
    sample_p(GA_ZP);
    GEACH(k, j) {
      const double qj = GAT(GA_Q, j), pj = GAT(GA_ZP, j), gj = GAT(GA_G0, j);
      GAT(GA_EQ0, j) = qj; GAT(GA_EQ1, j) = qj; GAT(GA_EP0, j) = pj; GAT(GA_EP1, j) = pj; GAT(GA_EG0, j) = gj; GAT(GA_EG1, j) = gj;
      GAT(GA_QS, j) = qj; GAT(GA_RHO, j) = pj;
    }
    lps = lp0; lsw = 0.0; H0 = -lp0 + kinetic(GA_ZP);
    depth = 0; n_leap = 0; sum_metro = 0.0; divergent = false;
    start_subtree();
  };
  auto begin_init_stepsize = [&](bool upd) {
    from_update = upd; teps = ad[0]; tdir = 0;
    if (teps == 0.0 || teps > 1e7) return false;
    prepare_trial();
    return true;
  };
  auto next_transition = [&]() {
    ++t;
    if (t >= a.n_steps) { phase = GPH_DONE; return; }
    rng.init(a.seed, (u64)(a.chain0 + cidx), a.step0 + (u64)t);
    GEACH(k, j) GAT(GA_QE, j) = GAT(GA_Q, j);
    phase = GPH_START;
  };
  auto end_transition = [&]() {
    const double accept_prob = sum_metro / (double)n_leap;
    GEACH(k, j) GAT(GA_Q, j) = GAT(GA_QS, j);
    lp = lps;
    ++n_acc;
    stat_depth = depth; stat_leaps += n_leap; stat_div += divergent ? 1.0 : 0.0; stat_acc = accept_prob;
    long long slot;
    if (active && record_slot(a, t, &slot)) GEACH(k, j) record_value(a, slot, j, chain, GAT(GA_Q, j));
    bool update = false;
    if (ad[9] == 1.0) {
      adapt_learn_stepsize(ad, cfg, accept_prob > 1.0 ? 1.0 : accept_prob);
      const bool in_w = adapt_in_window(ad, cfg), ends = adapt_window_ends(ad, cfg);
      if (in_w) { ad[8] += 1.0; GEACH(k, j) welford_add(ad[8], GAT(GA_Q, j), GAT(GA_WM, j), GAT(GA_WM2, j)); }
      if (ends) {
        adapt_next_window(ad, cfg);
        GEACH(k, j) { GAT(GA_MINV, j) = adapt_window_var(ad[8], GAT(GA_WM2, j), GAT(GA_MINV, j)); GAT(GA_WM, j) = 0.0; GAT(GA_WM2, j) = 0.0; }
        ad[8] = 0.0;
        update = true;
      }
      ad[5] += 1.0;
    }
    if (update) { GEACH(k, j) GAT(GA_QE, j) = GAT(GA_Q, j); phase = GPH_UPD; }
    else next_transition();
  };

  for (;;) {
    const bool busy = phase != GPH_DONE;
    if (!__any_sync(0xffffffffu, busy)) break;
    const double lpe = ev();                       // warp-wide: reads QE of every chain of the warp, fills GE
    if (!busy) continue;
    if (phase == GPH_START || phase == GPH_UPD) {
      lp0 = lpe; lp = lpe;
      GEACH(k, j) GAT(GA_G0, j) = GAT(GA_GE, j);
      if (phase == GPH_START) {
        bool trials = false;
        if (ad[9] == 0.0) {
          GEACH(k, j) { GAT(GA_MINV, j) = 1.0; GAT(GA_WM, j) = 0.0; GAT(GA_WM2, j) = 0.0; }
          adapt_reset(ad, cfg);
          trials = begin_init_stepsize(false);
        }
        if (!trials) begin_transition();
      } else if (!begin_init_stepsize(true)) {
        ad[1] = log(10.0 * ad[0]); ad[2] = 0.0; ad[3] = 0.0; ad[4] = 0.0;
        next_transition();
      }
    } else if (phase == GPH_TRIAL) {
      GEACH(k, j) GAT(GA_ZP, j) = fma(0.5 * teps, GAT(GA_GE, j), GAT(GA_ZP, j));
      double h = -lpe + kinetic(GA_ZP);
      if (isnan(h)) h = pos_inf();
      if (!init_stepsize_update(teps, tdir, H0t - h)) prepare_trial();
      else {
        ad[0] = teps;
        if (from_update) { ad[1] = log(10.0 * ad[0]); ad[2] = 0.0; ad[3] = 0.0; ad[4] = 0.0; next_transition(); }
        else begin_transition();
      }
    } else {  // GPH_LEAF
      const double e = sgn * eps;
      GEACH(k, j) { GAT(GA_ZQ, j) = GAT(GA_QE, j); GAT(GA_ZG, j) = GAT(GA_GE, j); GAT(GA_ZP, j) = fma(0.5 * e, GAT(GA_GE, j), GAT(GA_ZP, j)); }
      ++n_leap;
      double h = -lpe + kinetic(GA_ZP);
      if (isnan(h)) h = pos_inf();
      if (h - H0 > 1000.0) divergent = true;
      double c_lsw = H0 - h, c_lp = lpe;
      sum_metro += (H0 - h > 0.0) ? 1.0 : exp(H0 - h);
      GEACH(k, j) { GAT(GA_CRHO, j) = GAT(GA_ZP, j); GAT(GA_CPSL, j) = GAT(GA_MINV, j) * GAT(GA_ZP, j); GAT(GA_CQ, j) = GAT(GA_ZQ, j); }
      bool valid = !divergent;
      int kk = 0;
      while (valid && ((leaf >> kk) & 1)) {
        const double l_lsw = st_lsw[kk];
        const double sub = gen_log_sum_exp2(l_lsw, c_lsw);
        bool take_right = c_lsw > sub;
        if (!take_right) take_right = rng.u01() < exp(c_lsw - sub);
        double dl = 0.0, dr = 0.0;
        GEACH(k, j) {
          const double r = st_rho[kk][k] + GAT(GA_CRHO, j);
          GAT(GA_CRHO, j) = r;
          GAT(GA_CPSL, j) = st_psl[kk][k];
          if (!take_right) GAT(GA_CQ, j) = st_q[kk][k];
          dr = fma(GAT(GA_MINV, j) * GAT(GA_ZP, j), r, dr); dl = fma(GAT(GA_CPSL, j), r, dl);
        }
        dr = gsum(dr); dl = gsum(dl);
        if (!take_right) c_lp = st_lp[kk];
        c_lsw = sub;
        valid = dr > 0.0 && dl > 0.0;
        ++kk;
      }
      if (!valid) end_transition();
      else if (leaf + 1 == (1 << depth)) {
        const int eq = sgn > 0 ? GA_EQ1 : GA_EQ0, ep = sgn > 0 ? GA_EP1 : GA_EP0, eg = sgn > 0 ? GA_EG1 : GA_EG0;
        GEACH(k, j) { GAT(eq, j) = GAT(GA_ZQ, j); GAT(ep, j) = GAT(GA_ZP, j); GAT(eg, j) = GAT(GA_ZG, j); }
        ++depth;
        bool take = c_lsw > lsw;
        if (!take) take = rng.u01() < exp(c_lsw - lsw);
        lsw = gen_log_sum_exp2(lsw, c_lsw);
        double dp = 0.0, dm = 0.0;
        GEACH(k, j) {
          if (take) GAT(GA_QS, j) = GAT(GA_CQ, j);
          const double r = GAT(GA_RHO, j) + GAT(GA_CRHO, j);
          GAT(GA_RHO, j) = r;
          dp = fma(GAT(GA_MINV, j) * GAT(GA_EP1, j), r, dp); dm = fma(GAT(GA_MINV, j) * GAT(GA_EP0, j), r, dm);
        }
        dp = gsum(dp); dm = gsum(dm);
        if (take) lps = c_lp;
        if (!(dp > 0.0 && dm > 0.0) || depth >= max_depth) end_transition();
        else start_subtree();
      } else {
        GEACH(k, j) { st_rho[kk][k] = GAT(GA_CRHO, j); st_psl[kk][k] = GAT(GA_CPSL, j); st_q[kk][k] = GAT(GA_CQ, j); }
        st_lsw[kk] = c_lsw; st_lp[kk] = c_lp;
        ++leaf;
        prepare_leaf();
      }
    }
  }
  if (active) {
    GEACH(k, j) { f.Q[j * nc + chain] = GAT(GA_Q, j); f.MINV[j * nc + chain] = GAT(GA_MINV, j); f.WM[j * nc + chain] = GAT(GA_WM, j); f.WM2[j * nc + chain] = GAT(GA_WM2, j); }
  }
  if (writer) {
    for (int j = 0; j < STK_AD_LEN; ++j) f.AD[j * nc + chain] = ad[j];
    f.NS[0 * nc + chain] = stat_depth; f.NS[1 * nc + chain] += stat_leaps; f.NS[2 * nc + chain] += stat_div; f.NS[3 * nc + chain] = stat_acc;
    f.LP[chain] = lp;
    if (a.accept) a.accept[chain] += n_acc;
  }
#undef GAT
#undef GOWN
#undef GEACH
}

// ---- adapt_diag_e_static_hmc with the chain's state distributed over its lanes (the same layout and ownership as gen_nuts_lanes; arrays
// GA_Q, GA_MINV, GA_WM, GA_WM2, GA_QE = evaluation point, GA_GE = its gradient, GA_ZP = momenta, GA_G0 = gradient at the start of a
// trial, GA_QS = start of the trajectory).  All chains of a warp run the same loops to the warp's maximum trip count around the warp-wide
// evaluation, as gen_hmc_adapt does around a CTA-wide one.
template <int P, int G, int CPB, class Eval>
STK_DEV void gen_hmc_adapt_lanes(const StkArgs& a, const GenFields& f, const double* par_adapt, long long chain, long long cidx, bool active,
                                 int slice, int c, double* sm, Eval&& ev) {
  constexpr int NPAIR = (P + 1) / 2, NO = 2 * ((NPAIR + G - 1) / G);
  const long long nc = a.n_chains;
  const unsigned gmask = (G == 32 ? 0xffffffffu : ((1u << G) - 1u)) << ((threadIdx.x / G) * G);
  const bool writer = slice == 0 && active;
  AdaptCfg cfg; cfg.init(par_adapt);
#define GAT(arr, j) sm[((arr) * P + (j)) * CPB + c]
#define GOWN(k) (2 * (slice + ((k) >> 1) * G) + ((k) & 1))
#define GEACH(k, j) for (int k = 0; k < NO; ++k) if (const int j = GOWN(k); j < P)
  auto gsum = [&](double v) { return group_sum_masked<G>(v, gmask); };
  auto any = [&](bool b) { return (bool)__any_sync(0xffffffffu, b); };
  double ad[STK_AD_LEN];
  GEACH(k, j) { GAT(GA_Q, j) = f.Q[j * nc + cidx]; GAT(GA_MINV, j) = f.MINV[j * nc + cidx]; GAT(GA_QE, j) = GAT(GA_Q, j); }
  for (int j = 0; j < STK_AD_LEN; ++j) ad[j] = f.AD[j * nc + cidx];
  unsigned long long n_acc = 0;
  double lp = 0.0;
  auto sample_p = [&](Rng& rng) {
    const u32 s0 = rng.slot;
    for (int kk = 0; kk < NO / 2; ++kk) {
      const int b = slice + kk * G, j = 2 * b;
      if (j < P) {
        rng.slot = s0 + (u32)b;
        double z0, z1; rng.normal2(z0, z1);
        GAT(GA_ZP, j) = z0 / sqrt(GAT(GA_MINV, j)); if (j + 1 < P) GAT(GA_ZP, j + 1) = z1 / sqrt(GAT(GA_MINV, j + 1));
      }
    }
    rng.slot = s0 + (u32)NPAIR;
  };
  auto kinetic = [&]() { double ke = 0.0; GEACH(k, j) ke = fma(GAT(GA_MINV, j) * GAT(GA_ZP, j), GAT(GA_ZP, j), ke); return 0.5 * gsum(ke); };
  // evaluate at GA_Q (the evaluation point array is the mirror the evaluation reads)
  auto eval_at_q = [&]() { GEACH(k, j) GAT(GA_QE, j) = GAT(GA_Q, j); return ev(); };
  // base_hmc::init_stepsize; lp / GA_GE hold the density and gradient at q on entry and on exit
  auto init_stepsize = [&](Rng& rng, bool want) {
    double eps = ad[0];
    int dir = 0;
    bool done = !want || eps == 0.0 || eps > 1e7;
    const double lp0 = lp;
    GEACH(k, j) GAT(GA_G0, j) = GAT(GA_GE, j);
    while (any(!done)) {
      if (!done) sample_p(rng);
      const double H0 = -lp0 + kinetic();
      GEACH(k, j) { GAT(GA_ZP, j) = fma(0.5 * eps, GAT(GA_G0, j), GAT(GA_ZP, j)); GAT(GA_QE, j) = fma(eps * GAT(GA_MINV, j), GAT(GA_ZP, j), GAT(GA_Q, j)); }
      const double lpt = ev();
      GEACH(k, j) GAT(GA_ZP, j) = fma(0.5 * eps, GAT(GA_GE, j), GAT(GA_ZP, j));
      double h = -lpt + kinetic();
      if (isnan(h)) h = pos_inf();
      if (!done) done = init_stepsize_update(eps, dir, H0 - h);
    }
    if (want) ad[0] = eps;
    GEACH(k, j) GAT(GA_GE, j) = GAT(GA_G0, j);
  };
  for (long long t = 0; t < a.n_steps; ++t) {
    const unsigned long long step = a.step0 + (u64)t;
    Rng rng; rng.init(a.seed, (u64)(a.chain0 + cidx), step);
    const bool fresh = ad[9] == 0.0;
    if (fresh) { GEACH(k, j) GAT(GA_MINV, j) = 1.0; adapt_reset(ad, cfg); }
    lp = eval_at_q();
    if (any(fresh)) init_stepsize(rng, fresh);
    if (ad[9] == 1.0 && (double)step >= cfg.num_warmup) { ad[0] = exp(ad[3]); ad[9] = 2.0; }  // complete_adaptation
    const double eps = ad[0];
    const int L = adapt_leaps(cfg.int_time, eps);
    sample_p(rng);
    const double H0 = -lp + kinetic();
    const double lp0 = lp;
    GEACH(k, j) GAT(GA_QS, j) = GAT(GA_Q, j);
    for (int l = 0; any(l < L); ++l) {
      const bool on = l < L;
      if (on) GEACH(k, j) { GAT(GA_ZP, j) = fma(0.5 * eps, GAT(GA_GE, j), GAT(GA_ZP, j)); GAT(GA_Q, j) = fma(eps * GAT(GA_MINV, j), GAT(GA_ZP, j), GAT(GA_Q, j)); }
      lp = eval_at_q();
      if (on) GEACH(k, j) GAT(GA_ZP, j) = fma(0.5 * eps, GAT(GA_GE, j), GAT(GA_ZP, j));
    }
    double H = -lp + kinetic();
    if (isnan(H)) H = pos_inf();
    double accept_prob = exp(H0 - H);
    const double u = rng.u01();
    if (accept_prob < 1.0 && u > accept_prob) { GEACH(k, j) GAT(GA_Q, j) = GAT(GA_QS, j); lp = lp0; } else ++n_acc;
    if (accept_prob > 1.0) accept_prob = 1.0;
    bool update = false;
    if (ad[9] == 1.0) {
      adapt_learn_stepsize(ad, cfg, accept_prob);
      const bool in_w = adapt_in_window(ad, cfg), ends = adapt_window_ends(ad, cfg);
      if (in_w || ends) {
        // the Welford estimator stays in the module's wm / wm2 fields: every coordinate is read and written by its owner only
        GEACH(k, j) {
          double wm = f.WM[j * nc + cidx], wm2 = f.WM2[j * nc + cidx];
          if (in_w) welford_add(ad[8] + 1.0, GAT(GA_Q, j), wm, wm2);
          if (ends) { GAT(GA_MINV, j) = adapt_window_var(ad[8] + (in_w ? 1.0 : 0.0), wm2, GAT(GA_MINV, j)); wm = 0.0; wm2 = 0.0; }
          if (active) { f.WM[j * nc + chain] = wm; f.WM2[j * nc + chain] = wm2; }
        }
        if (in_w) ad[8] += 1.0;
        if (ends) { adapt_next_window(ad, cfg); ad[8] = 0.0; update = true; }
      }
      ad[5] += 1.0;
    }
    if (any(update)) {
      lp = eval_at_q();
      init_stepsize(rng, update);
      if (update) { ad[1] = log(10.0 * ad[0]); ad[2] = 0.0; ad[3] = 0.0; ad[4] = 0.0; }
    }
    long long slot;
    if (active && record_slot(a, t, &slot)) GEACH(k, j) record_value(a, slot, j, chain, GAT(GA_Q, j));
  }
  if (active) GEACH(k, j) { f.Q[j * nc + chain] = GAT(GA_Q, j); f.MINV[j * nc + chain] = GAT(GA_MINV, j); }
  if (writer) {
    for (int j = 0; j < STK_AD_LEN; ++j) f.AD[j * nc + chain] = ad[j];
    f.LP[chain] = lp;
    if (a.accept) a.accept[chain] += n_acc;
  }
#undef GAT
#undef GOWN
#undef GEACH
}

}  // namespace stk
#endif
