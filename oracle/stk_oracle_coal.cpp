// stk_oracle_coal.cpp — CPU ORACLE (test infrastructure) for config C2: the
// coal-mining reversible-jump change-point sampler
//     coal t `rjmc` coalMove' t          stochaskell/app/rj/Main.hs:115-238
// restated from the kernel IR that `runStep` prints, preserved in
// rj-coal.ipynb cell 8 (node names v_0_N below refer to that dump), and from
//     rjmc / rjmcRatio / rjmcTransRatio'      Data/Program.hs:1278-1347
//     pdfOrderStats                           Data/Program.hs:893-917
//     poissonProcess_lpdf                     Data/Expression/Eval.hs:185-190
//     lpdfs                                   Random/Distribution/Abstract.hs
//
// Parity status: pinned to the reference's own PRINTED PROGRAM, not to an executable reference — the reference can only run
// this kernel through the Haskell interpreter (CC and Stan lack pmf_lpdf / poissonProcess_lpdf / uniform_cdf, SURVEY 8a'),
// and GHC is absent.  The pin: the kernel IR that `runStep` printed (rj-coal.ipynb cell 8, tests/golden/ir/coal_rj.txt)
// is lowered node by node by the code generator (csrc/stkc.cpp) and run on B200; this restatement must agree with that
// generated kernel transition by transition under the generated kernel's own draws (orc_coal_rj_step_ir): dimension
// trajectory exactly, states 1e-9, log acceptance ratio / both densities / proposal ratio / log|det J| 1e-9
// (tests/test_coal_ir_gpu.py).  Further pins: the lpdf formulas (scipy / Boost elsewhere), the closed-form log|det J| per
// move against a finite-difference Jacobian of the move maps (tests/test_coal_cpu.py) and against the LU of the IR's block
// Jacobian on the device, and the statistical pins of rj-coal.ipynb (posterior mode of n in {2,3}).
//
// The only deviation from the IR: the IR evaluates log_det of the symbolic
// block Jacobian numerically (LU of a (2n+1+|y|+aux)^2 matrix, v_0_446);
// here each case uses the determinant's closed form (derivation in DESIGN.md).
#include "stk_oracle.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace {
const double NEG_INF = -std::numeric_limits<double>::infinity();

// 1-based: first j with e < v[j], else n+1   (CC.hs:192-197)
int find_sorted_insert_index(double e, const double* v, int n) {
  for (int i = 1; i <= n; ++i) if (e < v[i - 1]) return i;
  return n + 1;
}

// coalMoveProbs (app/rj/Main.hs:196-205; IR nodes v_0_0 .. v_0_21), arithmetic in the IR's order
void move_probs(int n, int lam, int nmax, double pr[4]) {
  const double c = (0.9 * (1.0 + (double)lam)) / (1.0 + 2.0 * (double)lam);                  // v_0_6
  const double b = (n == nmax) ? 0.0 : ((n < lam - 1) ? c : (c * (double)lam) / (1.0 + (double)n));   // v_0_11
  const double d = (n == 1) ? 0.0 : ((lam < n) ? c : (c * (double)n) / (double)lam);          // v_0_17
  const double r = (1.0 - b) - d;                                                             // v_0_18
  pr[0] = (n == 1) ? r : r / 2.0;    // e: rate move     v_0_20
  pr[1] = (n == 1) ? 0.0 : r / 2.0;  // p: position move v_0_21
  pr[2] = b;                         // birth
  pr[3] = d;                         // death
}

// lpdfAux' False (coal t) at (n,s,g,y): IR nodes v_0_64 .. v_0_97
double coal_logp(int n, const double* s, const double* g, const double* y, int ny, double t, int lam, double a, double b) {
  double lp_n = orc_poisson_lpdf(n, (double)lam);                                             // v_0_64 (no truncation constant)
  double lp_s = 0.0;
  if (!(1 > n)) {                                                                             // v_0_65
    // pdfOrderStats: every 2nd of 2n+1 ordered uniforms is observed (gaps of 1)
    double acc = orc_log_factorial(2 * n + 1);                                                // v_0_68
    acc += std::log(orc_uniform_cdf(s[0], 0.0, t));                                           // v_0_70 (x1 - logFactorial 1 folded away)
    double mid = 0.0;
    for (int i = 1; i <= n - 1; ++i) {                                                        // v_0_72, v_0_73
      double d = orc_uniform_cdf(s[i], 0.0, t) - orc_uniform_cdf(s[i - 1], 0.0, t);
      int gap = 2 * (i + 1) - 2 * i - 1;
      mid = (std::log(d) * (double)gap - orc_log_factorial(gap)) + mid;
    }
    acc += mid;
    int gr = (2 * n + 1) - 2 * n;
    acc += std::log(1.0 - orc_uniform_cdf(s[n - 1], 0.0, t)) * (double)gr - orc_log_factorial(gr);  // v_0_80
    double pts = 0.0;
    for (int i = 1; i <= n; ++i) pts = orc_uniform_lpdf(s[i - 1], 0.0, t) + pts;               // v_0_82
    acc += pts;
    lp_s = acc;                                                                               // v_0_83
  }
  double lp_g = 0.0;
  for (int i = 1; i <= n + 1; ++i) lp_g = lp_g + orc_gamma_lpdf(g[i - 1], a, b);               // v_0_87
  // coalLikelihood: area = sum_j w_j g_j, w = [s1, s_{i+1}-s_i ..., t - s_n]                 v_0_91 .. v_0_95
  double area = 0.0;
  for (int j = 1; j <= n + 1; ++j) {
    double w = (j == 1) ? s[0] : (j == n + 1 ? t - s[n - 1] : s[j - 1] - s[j - 2]);
    if (n == 0) w = t;
    area = w * g[j - 1] + area;
  }
  // poissonProcess_lpdf y t rate area = sum_i log rate(y_i) - area    (Eval.hs:185-190)
  double sl = 0.0;
  for (int i = 0; i < ny; ++i) sl += std::log(g[find_sorted_insert_index(y[i], s, n) - 1]);
  double lp_y = sl - area;                                                                    // v_0_96
  return ((lp_n + lp_s) + lp_g) + lp_y;                                                       // v_0_97 (+s, left fold)
}

bool vec_eq(const double* a, int na, const double* b, int nb) {
  if (na != nb) return false;
  for (int i = 0; i < na; ++i) if (a[i] != b[i]) return false;
  return true;
}
// foldr dflt [1..len] (\i r -> if a!i /= b!i then i else r): FIRST differing 1-based index (v_0_172, v_0_204 ...)
int first_diff(const double* a, const double* b, int len, int dflt, int na, int nb) {
  for (int i = 1; i <= len; ++i) {
    double av = i <= na ? a[i - 1] : std::numeric_limits<double>::quiet_NaN();  // out-of-range read: treated as "differs"
    double bv = i <= nb ? b[i - 1] : std::numeric_limits<double>::quiet_NaN();
    if (av != bv) return i;
  }
  return dflt;
}
double bound_s(const double* s, int n, int i, double t) { return i == 0 ? 0.0 : (i == n + 1 ? t : s[i - 1]); }

struct Trans { double lu_rev, lu_fwd, logdet; int kind; };

// rjmcTransRatio' (Program.hs:1306-1347) for coalMove': IR nodes v_0_131 .. v_0_446.
// x = (n,s,g) current, y = (n2,s2,g2) proposed.  Which move produced y is recovered by COMPARING the
// two states, exactly as the IR's guard conditions do.
Trans trans_ratio(int n, const double* s, const double* g, int n2, const double* s2, const double* g2,
                  double t, int lam, int nmax) {
  const bool s_eq = vec_eq(s2, n2, s, n);            // v_0_131
  const bool g_eq = vec_eq(g2, n2 + 1, g, n + 1);    // v_0_132
  const bool up = (n2 - n) == 1, down = (n - n2) == 1;
  const bool cH = s_eq && !g_eq && !up && !down;     // v_0_140
  const bool cP = g_eq && !up && !down && !s_eq;     // v_0_143
  const bool cB = !g_eq && up && !down && !s_eq;     // v_0_145
  const bool cD = !g_eq && !up && down && !s_eq;     // v_0_147
  Trans r; r.lu_rev = r.lu_fwd = r.logdet = 0.0; r.kind = 0;
  double px[4], py[4];
  move_probs(n, lam, nmax, px);
  move_probs(n2, lam, nmax, py);
  auto cat = [](const double* p, int k) { return std::log(p[k - 1]); };   // categorical_lpdf = pmf_lpdf, Const.hs:94
  if (cH) {
    r.kind = 1;
    int j = first_diff(g2, g, n2 + 1, -1, n2 + 1, n + 1);                                       // v_0_172 / v_0_227
    r.lu_rev = cat(py, 1) + orc_discrete_uniform_lpdf(j, 1, n2 + 1) +
               orc_uniform_lpdf(std::log(g[j - 1]) - std::log(g2[j - 1]), -0.5, 0.5);         // v_0_167,173,178
    r.lu_fwd = cat(px, 1) + orc_discrete_uniform_lpdf(j, 1, n + 1) +
               orc_uniform_lpdf(std::log(g2[j - 1]) - std::log(g[j - 1]), -0.5, 0.5);         // v_0_222,228,233
    r.logdet = std::log(g2[j - 1] / g[j - 1]);        // |det| = g'_j / g_j
  } else if (cP) {
    r.kind = 2;
    int j = first_diff(s2, s, n2, -1, n2, n);                                                   // v_0_182 / v_0_235
    double lo2 = bound_s(s2, n2, j - 1, t), hi2 = bound_s(s2, n2, j + 1, t);
    double lo = bound_s(s, n, j - 1, t), hi = bound_s(s, n, j + 1, t);
    r.lu_rev = cat(py, 2) + orc_discrete_uniform_lpdf(j, 1, n2) +
               orc_uniform_lpdf((s[j - 1] - lo2) / (hi2 - lo2), 0.0, 1.0);                    // v_0_168,183,201
    r.lu_fwd = cat(px, 2) + orc_discrete_uniform_lpdf(j, 1, n) +
               orc_uniform_lpdf((s2[j - 1] - lo) / (hi - lo), 0.0, 1.0);                      // v_0_223,236,254
    r.logdet = 0.0;                                   // |det| = 1
  } else if (cB) {
    r.kind = 3;
    int i = first_diff(s, s2, n, n + 1, n, n2);                                                 // v_0_218
    double xnew = s2[i - 1];
    int ig = first_diff(g, g2, n + 1, n + 2, n + 1, n2 + 1);                                    // v_0_259
    int j = find_sorted_insert_index(xnew, s, n);                                               // v_0_260
    double u = g2[ig - 1] / g[j - 1];                                                           // v_0_261
    r.lu_rev = cat(py, 4) + orc_discrete_uniform_lpdf(i, 1, n2);                                // v_0_169, v_0_219
    r.lu_fwd = cat(px, 3) + orc_uniform_lpdf(xnew, 0.0, t) + orc_normal_lpdf(std::log(u), 0.0, 1.0);  // v_0_225,256,263
    r.logdet = std::log(u * g[j - 1]);                // |det| = d(g'_{j+1})/dz = u g_j   (v_0_442)
  } else if (cD) {
    r.kind = 4;
    int i = first_diff(s2, s, n2, n2 + 1, n2, n);                                               // v_0_204
    double xold = s[i - 1];
    int ig = first_diff(g2, g, n2 + 1, n2 + 2, n2 + 1, n + 1);                                  // v_0_208
    int j = find_sorted_insert_index(xold, s2, n2);                                             // v_0_209
    double u = g[ig - 1] / g2[j - 1];                                                           // v_0_210
    r.lu_rev = cat(py, 3) + orc_uniform_lpdf(xold, 0.0, t) + orc_normal_lpdf(std::log(u), 0.0, 1.0);  // v_0_170,205,212
    r.lu_fwd = cat(px, 4) + orc_discrete_uniform_lpdf(i, 1, n);                                 // v_0_224, v_0_266
    r.logdet = -std::log(g[ig - 1]);                  // |det| = 1 / g_{j+1}      (v_0_419)
  }
  return r;
}

}  // namespace

// ---------------------------------------------------------------------------
// exported entry points
// ---------------------------------------------------------------------------
extern "C" double orc_coal_logp(int n, const double* s, const double* g, const double* y, int ny,
                                double t, int lam, double a, double b) {
  return coal_logp(n, s, g, y, ny, t, lam, a, b);
}
extern "C" void orc_coal_move_probs(int n, int lam, int nmax, double* pr4) { move_probs(n, lam, nmax, pr4); }

// log of the proposal/Jacobian part of the acceptance ratio (lu' - lu + logdet) and its three pieces
extern "C" int orc_coal_trans_ratio(int n, const double* s, const double* g, int n2, const double* s2, const double* g2,
                                    double t, int lam, int nmax, double* out3) {
  Trans r = trans_ratio(n, s, g, n2, s2, g2, t, lam, nmax);
  out3[0] = r.lu_rev; out3[1] = r.lu_fwd; out3[2] = r.logdet;
  return r.kind;
}

// The proposal `coalMove' t` (app/rj/Main.hs:207-214 = mixture' of coalMoveH/P/Birth'/Death').
// Philox consumption per transition (same layout in models/coal.cu):
//   block 0 -> (uA, uB): uA picks the move (pmf draw), uB is the accept uniform
//   block 1 -> (uC, uD): uC picks the index j (discreteUniform), uD is the continuous uniform (u or x)
//   block 2 -> z ~ normal(0,1) for the birth move (Box-Muller, cosine branch)
// The reference draws the auxiliaries of ALL four moves every transition (mixture' lowers to a pmf draw
// plus a `case`, IR bindings x_show_0_0..7); only the chosen move's are used, so drawing them lazily
// leaves the transition kernel's law unchanged.
// info[8] = {move, j, accepted, log ratio, f(y), u or x, z, f(x)}
// the deterministic part of one transition, given the auxiliary draws of the chosen move and the accept uniform:
// proposal (v_0_32 .. v_0_63), both densities (v_0_97, v_0_129), proposal/Jacobian ratio (v_0_131 .. v_0_446), decision
static int rj_apply(int* n_io, double* s, double* g, const double* y, int ny, double t, int lam, int nmax, double a, double b,
                    int move, int j, double aux, double z, double u_accept, double* info) {
  const int n = *n_io;
  int n2 = n;
  std::vector<double> s2(s, s + n), g2(g, g + n + 1);
  if (move == 1) {            // coalMoveH: g_j' = exp(log g_j + u)
    g2[j - 1] = std::exp(aux + std::log(g[j - 1]));                                             // v_0_33, v_0_34
  } else if (move == 2) {     // coalMoveP: s_j' = s(j-1) + (s(j+1) - s(j-1)) u
    double lo = bound_s(s, n, j - 1, t), hi = bound_s(s, n, j + 1, t);
    s2[j - 1] = lo + (hi - lo) * aux;                                                           // v_0_49 .. v_0_51
  } else if (move == 3) {     // coalMoveBirth': x ~ U(0,t), u = exp z, insert
    j = find_sorted_insert_index(aux, s, n);                                                    // v_0_53
    s2.insert(s2.begin() + (j - 1), aux);                                                       // v_0_54
    g2.insert(g2.begin() + j, std::exp(z) * g[j - 1]);                                          // v_0_56 .. v_0_58
    n2 = n + 1;
  } else {                    // coalMoveDeath'
    s2.erase(s2.begin() + (j - 1));                                                             // v_0_60
    g2.erase(g2.begin() + j);                                                                   // v_0_62
    n2 = n - 1;
  }
  const double fy = coal_logp(n2, s2.data(), g2.data(), y, ny, t, lam, a, b);                   // v_0_97
  const double fx = coal_logp(n, s, g, y, ny, t, lam, a, b);                                    // v_0_129
  Trans tr = trans_ratio(n, s, g, n2, s2.data(), g2.data(), t, lam, nmax);
  const double logr = ((fy - fx) + (tr.lu_rev - tr.lu_fwd)) + tr.logdet;                         // v_0_448
  const double ratio = std::fmin(1.0, std::exp(logr));                                           // v_0_450
  const int accept = u_accept < ratio ? 1 : 0;                                                   // bernoulli, CC.hs:290
  if (info) { info[0] = move; info[1] = j; info[2] = accept; info[3] = logr; info[4] = fy; info[5] = aux; info[6] = z; info[7] = fx; }
  if (accept) {
    *n_io = n2;
    std::memcpy(s, s2.data(), sizeof(double) * n2);
    std::memcpy(g, g2.data(), sizeof(double) * (n2 + 1));
  }
  return accept;
}

extern "C" int orc_coal_rj_step(int* n_io, double* s, double* g, const double* y, int ny, double t, int lam,
                                int nmax, double a, double b, uint64_t seed, uint64_t chain, uint64_t step,
                                double* info) {
  const int n = *n_io;
  double uA, uB, uC, uD;
  orc_u01x2(seed, chain, step, 0, &uA, &uB);
  orc_u01x2(seed, chain, step, 1, &uC, &uD);
  double pr[4];
  move_probs(n, lam, nmax, pr);
  // pmf draw (stk_device.cuh draw_pmf): first k with uA * total < running sum
  double tot = 0.0; for (int i = 0; i < 4; ++i) tot += pr[i];
  double ua = uA * tot, acc = 0.0; int move = 4;
  for (int i = 0; i < 4; ++i) { acc += pr[i]; if (ua < acc) { move = i + 1; break; } }
  auto dunif = [&](int lo, int hi) { int m = hi - lo + 1; int k = (int)std::floor(uC * (double)m); if (k >= m) k = m - 1; return lo + k; };
  int j = 0; double aux = 0.0, z = 0.0;
  if (move == 1) { j = dunif(1, n + 1); aux = -0.5 + (0.5 - (-0.5)) * uD; }
  else if (move == 2) { j = dunif(1, n); aux = 0.0 + (1.0 - 0.0) * uD; }
  else if (move == 3) { aux = 0.0 + (t - 0.0) * uD; double z1, z2; orc_normal2(seed, chain, step, 2, &z1, &z2); z = z1; }
  else j = dunif(1, n);
  return rj_apply(n_io, s, g, y, ny, t, lam, nmax, a, b, move, j, aux, z, uB, info);
}

// The same transition with the draws of the GENERATED kernel (stkc lowering of the printed IR, tests/golden/ir/coal_rj.txt):
// `mixture'` draws the auxiliaries of all four moves every transition (PNodes x_show_0_0 .. x_show_0_7) and the accept
// bernoulli is x_show_0_8; PNode x_<ns>_0_<P> draws with Philox counter word c2 ^ (P << 12), blocks from 0 (csrc/stkc.cpp).
extern "C" int orc_coal_rj_step_ir(int* n_io, double* s, double* g, const double* y, int ny, double t, int lam,
                                   int nmax, double a, double b, uint64_t seed, uint64_t chain, uint64_t step,
                                   double* info) {
  const int n = *n_io;
  auto gen = [&](uint32_t P) { orc_rng r; orc_rng_init(&r, seed, chain, step); r.c2 ^= (P << 12); r.slot = 0; return r; };
  double pr[4];
  move_probs(n, lam, nmax, pr);
  orc_rng r0 = gen(0); const int move = orc_draw_pmf(&r0, pr, 4);                               // x_show_0_0
  orc_rng r1 = gen(1); const int j_rate = orc_draw_discrete_uniform(&r1, 1, n + 1);             // x_show_0_1
  orc_rng r2 = gen(2); const double u_rate = -0.5 + (0.5 - (-0.5)) * orc_u01(&r2);              // x_show_0_2
  orc_rng r3 = gen(3); const int j_pos = orc_draw_discrete_uniform(&r3, 1, n);                  // x_show_0_3
  orc_rng r4 = gen(4); const double u_pos = 0.0 + (1.0 - 0.0) * orc_u01(&r4);                   // x_show_0_4
  orc_rng r5 = gen(5); const double x_birth = 0.0 + (t - 0.0) * orc_u01(&r5);                   // x_show_0_5
  orc_rng r6 = gen(6); const double z_birth = 0.0 + 1.0 * orc_normal(&r6);                      // x_show_0_6
  orc_rng r7 = gen(7); const int j_death = orc_draw_discrete_uniform(&r7, 1, n);                // x_show_0_7
  orc_rng r8 = gen(8); const double u_acc = orc_u01(&r8);                                       // x_show_0_8 = (u < p)
  int j = move == 1 ? j_rate : move == 2 ? j_pos : move == 4 ? j_death : 0;
  double aux = move == 1 ? u_rate : move == 2 ? u_pos : move == 3 ? x_birth : 0.0;
  return rj_apply(n_io, s, g, y, ny, t, lam, nmax, a, b, move, j, aux, z_birth, u_acc, info);
}

// coalPrior (app/rj/Main.hs:115-122): n ~ truncated 1 nMax (poisson lam) by rejection (CC.hs:282-286),
// s = every 2nd of 2n+1 sorted U(0,t), g_j ~ gamma a b.  One sequential Philox stream, step = 2^64-1.
extern "C" void orc_coal_simulate(uint64_t seed, uint64_t chain, double t, int lam, int nmax, double a, double b,
                                  int* n_out, double* s, double* g) {
  orc_rng r; orc_rng_init(&r, seed, chain, ~0ULL);
  int n = 1;
  for (int it = 0; it < 1000; ++it) { n = orc_draw_poisson(&r, (double)lam); if (n >= 1 && n <= nmax) break; }
  if (n < 1) n = 1;
  if (n > nmax) n = nmax;
  std::vector<double> u(2 * n + 1);
  for (auto& v : u) v = 0.0 + (t - 0.0) * orc_u01(&r);
  std::sort(u.begin(), u.end());
  for (int i = 1; i <= n; ++i) s[i - 1] = u[2 * i - 1];     // s'!(2i), 1-based
  for (int j = 0; j < n + 1; ++j) g[j] = orc_draw_gamma(&r, a, b);
  *n_out = n;
}

// CPU-baseline driver: C chains x T transitions, one sequential chain per host thread at a time.
// State SoA as on the GPU: n[C] (int32), s[smax][C], g[smax+1][C] with smax = nmax.
extern "C" long long orc_coal_rj_run(int* n, double* s, double* g, long long C, long long T, const double* y, int ny,
                                     double t, int lam, int nmax, double a, double b, uint64_t seed,
                                     uint64_t chain0, uint64_t step0, int nthreads) {
  std::atomic<long long> next{0}, accepted{0};
  auto work = [&]() {
    std::vector<double> sc(nmax + 1), gc(nmax + 2);
    for (;;) {
      long long c = next.fetch_add(1);
      if (c >= C) break;
      int nc = n[c];
      for (int i = 0; i < nc; ++i) sc[i] = s[(size_t)i * C + c];
      for (int i = 0; i < nc + 1; ++i) gc[i] = g[(size_t)i * C + c];
      long long acc = 0;
      for (long long k = 0; k < T; ++k)
        acc += orc_coal_rj_step(&nc, sc.data(), gc.data(), y, ny, t, lam, nmax, a, b, seed, chain0 + c, step0 + k, nullptr);
      n[c] = nc;
      for (int i = 0; i < nmax; ++i) s[(size_t)i * C + c] = i < nc ? sc[i] : 0.0;
      for (int i = 0; i < nmax + 1; ++i) g[(size_t)i * C + c] = i < nc + 1 ? gc[i] : 0.0;
      accepted.fetch_add(acc);
    }
  };
  if (nthreads <= 1) { work(); return accepted.load(); }
  std::vector<std::thread> th;
  for (int i = 0; i < nthreads; ++i) th.emplace_back(work);
  for (auto& x : th) x.join();
  return accepted.load();
}
