// stk_oracle.h — CPU ORACLE (test infrastructure, NOT product code).
//
// A plain C++ restatement of the reference's algorithms for the sampling hot
// path, used only by tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs as the CHECKER.  The product path
// (libstk_cuda.so) never links, loads or calls anything in oracle/.
//
// Parity status: PINNED for the pieces noted below, "parity unpinned" for the
// rest:
//   * lpdf formulas: restated from Abstract.hs / Const.hs; pinned against the
//     vendored Boost.Math (the CC backend's own arithmetic, CC.hs:206-228) by
//     oracle/ref_cc/*.cc compiled from the reference headers, and against the
//     vendored Stan (lp + gradient via `diagnose`) through tests/golden/.
//   * Philox4x32-10: pinned by the Random123 known-answer vectors.
//   * soccer RJ kernel: pinned against the reference's own generated C++
//     (rj.ipynb cell 10) compiled into oracle/_ref/.
//   * coal RJ kernel: restated node by node from the kernel IR preserved in
//     rj-coal.ipynb cell 8; the reference can only run it through the Haskell
//     interpreter (no GHC here) -> statistical pins only.
#ifndef STK_ORACLE_H
#define STK_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

// ---- RNG: same counter layout as stk_device.cuh ------------------------------
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
// k-th block of (seed, chain, step): words -> two uniforms / two normals
void orc_u01x2(uint64_t seed, uint64_t chain, uint64_t step, uint32_t slot, double* a, double* b);
void orc_normal2(uint64_t seed, uint64_t chain, uint64_t step, uint32_t slot, double* a, double* b);
// stateful stream mirroring stk::Rng
typedef struct { uint32_t k0, k1, c0, c1, c2, slot; } orc_rng;
void orc_rng_init(orc_rng* g, uint64_t seed, uint64_t chain, uint64_t step);
double orc_u01(orc_rng* g);
double orc_normal(orc_rng* g);
void orc_rng_normal2(orc_rng* g, double* a, double* b);  /* one block -> two normals */
int orc_draw_discrete_uniform(orc_rng* g, int a, int b);
int orc_draw_bernoulli(orc_rng* g, double p);
int orc_draw_pmf(orc_rng* g, const double* ps, int n);
int orc_draw_poisson(orc_rng* g, double l);
double orc_draw_gamma(orc_rng* g, double a, double b);
double orc_draw_cauchy(orc_rng* g, double m, double s);
double orc_draw_inv_gamma(orc_rng* g, double a, double b);
int orc_draw_geometric(orc_rng* g, double p);
double orc_draw_logistic(orc_rng* g, double m, double s);
int orc_draw_neg_binomial(orc_rng* g, double a, double b);
double orc_draw_beta(orc_rng* g, double a, double b);
double orc_draw_lkj_corr2(orc_rng* g, double v);

// ---- log densities (Abstract.hs:68-167, Const.hs:90-106) ---------------------
double orc_normal_lpdf(double x, double m, double s);
double orc_gamma_lpdf(double x, double a, double b);
double orc_inv_gamma_lpdf(double x, double a, double b);
double orc_geometric_lpdf(int k, double p);
double orc_neg_binomial_lpdf(int k, double a, double b);
double orc_poisson_lpdf(int k, double l);
double orc_uniform_lpdf(double x, double a, double b);
double orc_uniform_cdf(double x, double a, double b);
double orc_discrete_uniform_lpdf(int x, int a, int b);
double orc_bernoulli_logit_lpdf(int b, double l);
double orc_cauchy_lpdf(double x, double m, double s);
double orc_log_factorial(int k);

// ---- C1 linreg (linreg.ipynb cell 1, d-dimensional) --------------------------
// q = (b, w[d], log sigma); X row-major [n][d].  Returns lp; grad[d+2] if non-NULL.
double orc_linreg_logp(const double* X, const double* y, int n, int d, const double* q, double* grad);
// one static-HMC transition (base_static_hmc.hpp:42-72) with the same Philox
// consumption as linreg_hmc; q updated in place; returns 1 if accepted.
int orc_linreg_hmc_step(const double* X, const double* y, int n, int d, double* q, double eps, int L,
                        uint64_t seed, uint64_t chain, uint64_t step, double* lp_out);
int orc_linreg_mh_step(const double* X, const double* y, int n, int d, double* q, double rw,
                       uint64_t seed, uint64_t chain, uint64_t step, double* lp_io);
// forward simulation: q[d+2] and ysim[n]
void orc_linreg_simulate(const double* X, int n, int d, uint64_t seed, uint64_t chain, double* q, double* ysim);
// CPU-baseline driver: C chains x T transitions on nthreads host threads; q is [P][C]; returns #accepted
long long orc_linreg_hmc_run(const double* X, const double* y, int n, int d, double* q, long long C, long long T,
                             double eps, int L, uint64_t seed, uint64_t chain0, uint64_t step0, int nthreads);

// ---- C2 coal-mining RJ (app/rj/Main.hs:115-238; rj-coal.ipynb cell 8) ---------
double orc_coal_logp(int n, const double* s, const double* g, const double* y, int ny, double t, int lam, double a, double b);
void orc_coal_move_probs(int n, int lam, int nmax, double* pr4);
int orc_coal_trans_ratio(int n, const double* s, const double* g, int n2, const double* s2, const double* g2,
                         double t, int lam, int nmax, double* out3);
int orc_coal_rj_step(int* n_io, double* s, double* g, const double* y, int ny, double t, int lam, int nmax,
                     double a, double b, uint64_t seed, uint64_t chain, uint64_t step, double* info8);
/* the same transition under the Philox layout of the stkc-generated kernel (9 PNodes of the printed IR) */
int orc_coal_rj_step_ir(int* n_io, double* s, double* g, const double* y, int ny, double t, int lam, int nmax,
                     double a, double b, uint64_t seed, uint64_t chain, uint64_t step, double* info8);
void orc_coal_simulate(uint64_t seed, uint64_t chain, double t, int lam, int nmax, double a, double b,
                       int* n_out, double* s, double* g);
long long orc_coal_rj_run(int* n, double* s, double* g, long long C, long long T, const double* y, int ny,
                          double t, int lam, int nmax, double a, double b, uint64_t seed, uint64_t chain0,
                          uint64_t step0, int nthreads);

// ---- C4 birats (birats.ipynb; q = betaMu[2], tau[2], y_omega, beta[.,1], beta[.,2], log sigma) ----
double orc_birats_logp(const double* x, const double* y, int N, int Tn, const double* q, double* grad);
int orc_birats_hmc_step(const double* x, const double* y, int N, int Tn, double* q, double eps, int L,
                        uint64_t seed, uint64_t chain, uint64_t step, double* lp_out);
long long orc_birats_hmc_run(const double* x, const double* y, int N, int Tn, double* q, long long C, long long T,
                             double eps, int L, uint64_t seed, uint64_t chain0, uint64_t step0, int nthreads);

// ---- C5 DP mixture (dp.ipynb; truncated stick-breaking, blocked Gibbs; PARITY UNPINNED) -------
void orc_dp_sweep(const double* y, int N, int K, double alpha, double lo, double hi, double* v, double* mu, int* z,
                  uint64_t seed, uint64_t chain, uint64_t sweep);
void orc_dp_simulate(int N, int K, double alpha, double lo, double hi, uint64_t seed, uint64_t chain,
                     double* v, double* mu, int* z, double* ysim);
long long orc_dp_run(const double* y, int N, int K, double alpha, double lo, double hi, double* v, double* mu,
                     long long C, long long T, uint64_t seed, uint64_t chain0, uint64_t step0, int nthreads);

// ---- C3 GP classifier, gp.ipynb form (fixed SE kernel, whitened latent w; pinned by tests/golden/gp_stan.npz) ----
int orc_gp_chol(const double* x, int n, double* L);
double orc_gp_logp(const double* L, const int* z, int n, const double* w, double* grad);
int orc_gp_hmc_step(const double* L, const int* z, int n, double* w, double eps, int Lf,
                    uint64_t seed, uint64_t chain, uint64_t step, double* lp_out);
long long orc_gp_hmc_run(const double* L, const int* z, int n, double* W, long long C, long long T, double eps, int Lf,
                         uint64_t seed, uint64_t chain0, uint64_t step0, int nthreads);

// ---- C3 with per-chain hyper-parameters: GP marginal likelihood, one Cholesky per evaluation (stk_oracle_gpr.cpp;
//      pinned by tests/golden/gpr_cc.json = the CC templates on the vendored Eigen) ----
double orc_gpr_logp(const double* x, const double* y, int n, const double* q);
int orc_gpr_mh_step(const double* x, const double* y, int n, double* q, double rw, uint64_t seed, uint64_t chain,
                    uint64_t step, double* lp_io);
long long orc_gpr_mh_run(const double* x, const double* y, int n, double* Q, double* LP, long long C, long long T, double rw,
                         uint64_t seed, uint64_t chain0, uint64_t step0, int nthreads);

// ---- soccer RJ (rj.ipynb; pinned against the reference's generated C++ in oracle/_ref) -----
double orc_soccer_accept_prob(int tag, double lam, double kappa, const int* y, int ny, double x, double* kappa_prop);
int orc_soccer_rj_step(int tag, double lam, double* kappa, const int* y, int ny,
                       uint64_t seed, uint64_t chain, uint64_t step, double* p_out);

// ---- adaptive static HMC with a diagonal metric (Stan's adapt_diag_e_static_hmc; stk_oracle_adapt.cpp) ----
#define ORC_AD_LEN 10
#define ORC_CFG_LEN 10
int orc_linreg_hmc_adapt_step(const double* X, const double* y, int n, int d, double* q, double* ad, double* minv,
                              double* wm, double* wm2, const double* cfg, uint64_t seed, uint64_t chain, uint64_t step,
                              double* lp_out, double* out3);
int orc_birats_hmc_adapt_step(const double* x, const double* y, int N, int Tn, double* q, double* ad, double* minv,
                              double* wm, double* wm2, const double* cfg, uint64_t seed, uint64_t chain, uint64_t step,
                              double* lp_out, double* out3);

// NUTS with a diagonal metric + the same adaptation (hmcStan's default sampler); out5: accept_stat, eps, treedepth, n_leapfrog, divergent
void orc_linreg_nuts_step(const double* X, const double* y, int n, int d, double* q, double* ad, double* minv, double* wm,
                          double* wm2, const double* cfg, int max_depth, uint64_t seed, uint64_t chain, uint64_t step,
                          double* lp_out, double* out5);
void orc_birats_nuts_step(const double* x, const double* y, int N, int Tn, double* q, double* ad, double* minv, double* wm,
                          double* wm2, const double* cfg, int max_depth, uint64_t seed, uint64_t chain, uint64_t step,
                          double* lp_out, double* out5);

long long orc_linreg_nuts_run(const double* X, const double* y, int n, int d, double* q, double* ad, double* minv, double* wm,
                              double* wm2, long long C, long long T, const double* cfg, int max_depth, uint64_t seed,
                              uint64_t chain0, uint64_t step0, int nthreads);
long long orc_birats_nuts_run(const double* x, const double* y, int N, int Tn, double* q, double* ad, double* minv, double* wm,
                              double* wm2, long long C, long long T, const double* cfg, int max_depth, uint64_t seed,
                              uint64_t chain0, uint64_t step0, int nthreads);

#ifdef __cplusplus
}
#endif
#endif
