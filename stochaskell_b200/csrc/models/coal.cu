// coal.cu — hand-tuned module for config C2: the coal-mining reversible-jump
// change-point sampler        coal t `rjmc` coalMove' t
//   model     stochaskell/app/rj/Main.hs:115-137  (coalPrior, coalLikelihood)
//   proposal  stochaskell/app/rj/Main.hs:139-214  (coalMoveH/P/Birth'/Death', coalMoveProbs, mixture')
//   kernel    rjmc / rjmcRatio / rjmcTransRatio'  Data/Program.hs:1278-1347
//   kernel IR rj-coal.ipynb cell 8 (456 top-level nodes; v_0_N in comments)
//
// Pattern ("variable-dimension state, G lanes per chain"):
//   * state (n, s[1..n], g[1..n+1]) of every chain of the CTA lives in SHARED
//     memory for the whole launch, 1-based with sentinels s[0]=0, s[n+1]=t, so
//     that `replaceIndex/insertIndex/deleteIndex` (CC.hs:169-191) are index
//     arithmetic on the read side: the proposed state is never materialised,
//     f(y) reads it through a shifted accessor; an accepted birth/death is an
//     in-place shift by the chain's G lanes.
//   * the joint log-density is a sum over the n+1 segments; lane l of the chain's
//     group evaluates segments l+1, l+1+G, ... and the group combines them with a
//     fixed-order shuffle butterfly (deterministic).
//   * sum_i log rate(y_i) (poissonProcess_lpdf, Eval.hs:185-190) over the 191
//     sorted event times is sum_j cnt_j log g_j with cnt_j from ONE binary search
//     per segment in the smem copy of y (instead of 191 sorted-insert searches).
//   * move probabilities / their logs / log-factorial heads are per-n tables
//     built once per launch in smem with the IR's own operation order.
//   * log|det J| of the transformation (v_0_446; the reference LU-factorises a
//     (2n+1+|y|+aux)^2 block matrix numerically) in closed form per move:
//       rate move  log(g_j'/g_j);  position move 0;
//       birth      log(g'_{j+1});  death  -log(g_{j+1}).
//   * f(x) of the current state is carried between transitions (the reference
//     re-evaluates it, v_0_98..v_0_129; same value).
// Philox consumption per transition (restated bit-for-bit in oracle/stk_oracle_coal.cpp):
//   block 0 -> (uA, uB): move type (pmf), accept uniform
//   block 1 -> (uC, uD): index j (discreteUniform), continuous uniform (u / x)
//   block 2 -> z ~ normal(0,1) for the birth move
#include "../stk_device.cuh"
#include "../stk_fastmath.cuh"

using namespace stk;

namespace {

constexpr int SMAX = 30;          // nMax (app/rj/Main.hs:117, 199)
constexpr int ROW = 32;           // s[0..31], g[0..31] per chain
constexpr int BLOCK = 128;

enum { F_N = 0, F_S = 1, F_G = 2, F_LP = 3, F_Y = 4 };
enum { PAR_T = 0, PAR_LAM = 1, PAR_NMAX = 2, PAR_A = 3, PAR_B = 4 };

// per-launch tables (index n = 0..31)
struct Tabs {
  double head[32];      // poisson_lpdf(n; lam) + logFactorial(2n+1) + n log(1/t)     v_0_64, v_0_68, v_0_82
  double logdinv[34];   // log(1 / m)  = discreteUniform_lpdf(.; 1, m)                 Abstract.hs:166-167
  double pr[32][4];     // coalMoveProbs n = (e, p, b, d)                              v_0_0 .. v_0_21
  double lpr[32][4];    // log of the above (pmf_lpdf, Const.hs:94)
};
constexpr int TAB_DOUBLES = sizeof(Tabs) / 8;

__device__ __forceinline__ void build_tabs(Tabs* T, double t, int lam, int nmax) {
  for (int n = threadIdx.x; n < 34; n += blockDim.x) {
    T->logdinv[n] = n > 0 ? log(1.0 / (double)n) : 0.0;
    if (n >= 32) continue;
    const double dl = (double)lam, dn = (double)n;
    T->head[n] = ((double)n * log(dl) - lgamma(1.0 + dn) - dl) + lgamma(1.0 + (double)(2 * n + 1)) + dn * log(1.0 / t);
    const double c = (0.9 * (1.0 + dl)) / (1.0 + 2.0 * dl);
    const double b = (n == nmax) ? 0.0 : ((n < lam - 1) ? c : (c * dl) / (1.0 + dn));
    const double d = (n == 1) ? 0.0 : ((lam < n) ? c : (c * dn) / dl);
    const double r = (1.0 - b) - d;
    const double pr[4] = {(n == 1) ? r : r / 2.0, (n == 1) ? 0.0 : r / 2.0, b, d};
#pragma unroll
    for (int k = 0; k < 4; ++k) { T->pr[n][k] = pr[k]; T->lpr[n][k] = log(pr[k]); }
  }
}

// number of y_i < v  (y ascending): the 0-based lower bound
__device__ __forceinline__ int count_below(const double* __restrict__ y, int ny, double v) {
  int lo = 0, len = ny;
  while (len > 0) {
    const int half = len >> 1;
    const bool lt = y[lo + half] < v;
    lo = lt ? lo + half + 1 : lo;
    len = lt ? len - half - 1 : half;
  }
  return lo;
}

// A proposed state seen through the current one.
//   move 0: none, 1: g_j := gnew, 2: s_j := snew, 3: insert snew at j / gnew at j+1, 4: delete s_j / g_{j+1}
struct View {
  const double* s; const double* g;
  int sgn, thr_s, thr_g, is, ig; double snew, gnew;
  __device__ __forceinline__ void set(const double* s_, const double* g_, int move, int j, double sn, double gn) {
    s = s_; g = g_; snew = sn; gnew = gn;
    sgn = move == 3 ? -1 : (move == 4 ? 1 : 0);
    thr_s = move == 3 ? j : j - 1;          // birth shifts i > j, death shifts i >= j
    thr_g = j;                              // both shift i > j
    is = (move == 2 || move == 3) ? j : -1;
    ig = move == 1 ? j : (move == 3 ? j + 1 : -1);
  }
  __device__ __forceinline__ double ps(int i) const { const double v = s[i + (i > thr_s ? sgn : 0)]; return i == is ? snew : v; }
  __device__ __forceinline__ double pg(int i) const { const double v = g[i + (i > thr_g ? sgn : 0)]; return i == ig ? gnew : v; }
};

struct Consts { double t, a, b, alb, am1; int ny; };

// joint log-density of the viewed state with np change points (lpdfAux' False (coal t), v_0_64..v_0_97);
// all G lanes of the group return the same value.  `rounds` is warp-uniform.
template <int G>
__device__ __forceinline__ double eval_f(const View& v, int np, const double* __restrict__ ysm, const Tabs* T,
                                         const Consts& k, int l, int rounds) {
  double acc = 0.0, carry_hi = 0.0, carry_F = 0.0;
  int carry_c = 0;
  for (int r = 0; r < rounds; ++r) {
    const int i = r * G + l + 1;
    const bool valid = i <= np + 1;
    const int ii = valid ? i : np + 1;
    const double hi = v.ps(ii), gi = v.pg(ii);
    const bool inner = ii <= np;
    const int c_hi = inner ? count_below(ysm, k.ny, hi) : k.ny;
    const double F = hi < 0.0 ? 0.0 : (hi > k.t ? 1.0 : hi / k.t);          // uniform_cdf, Abstract.hs:163-165
    double lo = __shfl_up_sync(0xffffffffu, hi, 1, G);
    double Flo = __shfl_up_sync(0xffffffffu, F, 1, G);
    int c_lo = __shfl_up_sync(0xffffffffu, c_hi, 1, G);
    if (l == 0) { lo = carry_hi; Flo = carry_F; c_lo = carry_c; }
    if (G < 32) {
      carry_hi = __shfl_sync(0xffffffffu, hi, G - 1, G);
      carry_F = __shfl_sync(0xffffffffu, F, G - 1, G);
      carry_c = __shfl_sync(0xffffffffu, c_hi, G - 1, G);
    }
    const double lg = log(gi);
    double term = log(F - Flo)                                   // order statistics: v_0_70, v_0_72, v_0_80
                  + ((k.alb + k.am1 * lg) - k.b * gi)            // gamma_lpdf g_i a b, Abstract.hs:68
                  + ((double)(c_hi - c_lo) * lg - (hi - lo) * gi);  // poissonProcess_lpdf share of segment i
    if (inner && (hi < 0.0 || k.t < hi)) term = neg_inf();       // uniform_lpdf s_i 0 t, Abstract.hs:161-162
    acc += valid ? term : 0.0;
  }
  return group_sum<G>(acc) + T->head[np];
}

}  // namespace

namespace {

template <int G>
struct Chain {
  static constexpr int CPB = BLOCK / G;
  double* ysm; Tabs* T; double* s; double* g;
  Consts k; int lam, nmax;
  int l, grp; long long chain, cidx; bool active;

  __device__ __forceinline__ void setup(const StkArgs& a, double* smem) {
    const int ny = (int)a.flen[F_Y];
    const int nyp = (ny + 1) & ~1;
    ysm = smem; T = reinterpret_cast<Tabs*>(smem + nyp);
    l = threadIdx.x % G; grp = threadIdx.x / G;
    s = smem + nyp + TAB_DOUBLES + grp * (2 * ROW); g = s + ROW;
    k.t = a.par[PAR_T]; k.a = a.par[PAR_A]; k.b = a.par[PAR_B]; k.ny = ny;
    k.alb = k.a * log(k.b) - lgamma(k.a); k.am1 = k.a - 1.0;
    lam = (int)a.par[PAR_LAM]; nmax = min((int)a.par[PAR_NMAX], SMAX);
    const double* Y = (const double*)a.field[F_Y];
    for (int i = threadIdx.x; i < ny; i += blockDim.x) ysm[i] = Y[i];
    build_tabs(T, k.t, lam, nmax);
    chain = (long long)blockIdx.x * CPB + grp;
    active = chain < a.n_chains;
    cidx = active ? chain : a.n_chains - 1;
  }
  // HBM -> smem, with the sentinels; returns n
  __device__ __forceinline__ int load(const StkArgs& a) {
    int n = ((const int*)a.field[F_N])[cidx];
    n = max(0, min(n, SMAX));
    const double* S = (const double*)a.field[F_S];
    const double* Gm = (const double*)a.field[F_G];
    for (int i = l + 1; i < ROW; i += G) {
      s[i] = (i <= n) ? S[(long long)(i - 1) * a.n_chains + cidx] : 0.0;
      g[i] = (i <= n + 1) ? Gm[(long long)(i - 1) * a.n_chains + cidx] : 0.0;
    }
    if (l == 0) { s[0] = 0.0; g[0] = 0.0; }
    __syncwarp();
    if (l == 0) s[n + 1] = k.t;
    __syncwarp();
    return n;
  }
  __device__ __forceinline__ void store(const StkArgs& a, int n, double lp) {
    if (!active) return;
    double* S = (double*)a.field[F_S];
    double* Gm = (double*)a.field[F_G];
    for (int i = l + 1; i < ROW; i += G) {
      if (i <= SMAX) S[(long long)(i - 1) * a.n_chains + chain] = (i <= n) ? s[i] : 0.0;
      if (i <= SMAX + 1) Gm[(long long)(i - 1) * a.n_chains + chain] = (i <= n + 1) ? g[i] : 0.0;
    }
    if (l == 0) { ((int*)a.field[F_N])[chain] = n; ((double*)a.field[F_LP])[chain] = lp; }
  }
  __device__ __forceinline__ int warp_rounds(int np) const {
    return __reduce_max_sync(0xffffffffu, (np + 1 + G - 1) / G);
  }
};

template <int G>
__device__ __forceinline__ void rj_body(const StkArgs& a) {
  extern __shared__ __align__(16) double smem[];
  Chain<G> ch; ch.setup(a, smem);
  __syncthreads();
  int n = ch.load(a);
  const Consts& k = ch.k; const Tabs* T = ch.T; double* s = ch.s; double* g = ch.g; const int l = ch.l;
  View v; v.set(s, g, 0, 0, 0.0, 0.0);
  double lp = eval_f<G>(v, n, ch.ysm, T, k, l, ch.warp_rounds(n));
  const double log_inv_t = log(1.0 / (k.t - 0.0));           // uniform_lpdf x 0 t, Abstract.hs:161-162
  const u64 gchain = (u64)(a.chain0 + ch.cidx);
  unsigned long long n_acc = 0;

  for (long long tr = 0; tr < a.n_steps; ++tr) {
    Rng rng; rng.init(a.seed, gchain, a.step0 + (u64)tr);
    double uA, uB, uC, uD;
    rng.u01x2(uA, uB);
    rng.u01x2(uC, uD);
    // x_show_0_0 <- pmf [e,p,b,d]   (draw_pmf)
    const double* pr = T->pr[n];
    const double tot = ((pr[0] + pr[1]) + pr[2]) + pr[3];
    const double ua = uA * tot;
    double cum = pr[0]; int move = 4;
    if (ua < cum) move = 1;
    else { cum += pr[1]; if (ua < cum) move = 2; else { cum += pr[2]; if (ua < cum) move = 3; } }
    // index draw: discreteUniform 1 (n+1) for the rate move, 1 n otherwise (unused by birth)
    const int m = max((move == 1) ? n + 1 : n, 1);
    int j = (int)floor(uC * (double)m); if (j >= m) j = m - 1; j += 1;
    // birth: x ~ uniform 0 t, j = findSortedInsertIndex x s  (v_0_53) by ballot over the group
    const double x = 0.0 + (k.t - 0.0) * uD;
    int jb = 1;
    {
      const int rounds = ch.warp_rounds(n - 1);          // ceil(n / G), warp-uniform
      for (int r = 0; r < rounds; ++r) {
        const int i = r * G + l + 1;
        const bool le = (i <= n) && !(x < s[i]);
        unsigned b = __ballot_sync(0xffffffffu, le);
        if (G < 32) b = (b >> ((threadIdx.x & 31) / G * G)) & ((1u << G) - 1u);
        jb += __popc(b);
      }
    }
    int np = n; double snew = 0.0, gnew = 0.0, lq = 0.0;   // lq = (lu' - lu) + log|det J|
    if (move == 1) {                       // coalMoveH: g_j' = exp(log g_j + u), u ~ uniform -1/2 1/2
      const double u = -0.5 + (0.5 - (-0.5)) * uD;
      const double gj = g[j];
      gnew = exp(u + log(gj));
      lq = log(gnew / gj);               // lu' = lu (same n, same j, uniform densities equal); |det J| = g_j'/g_j
    } else if (move == 2) {                // coalMoveP: s_j' = s(j-1) + (s(j+1) - s(j-1)) u, u ~ uniform 0 1
      const double lo = s[j - 1], hi = s[j + 1];
      snew = lo + (hi - lo) * (0.0 + (1.0 - 0.0) * uD);
      lq = 0.0;
    } else if (move == 3) {                // coalMoveBirth': insert x at j, g'_{j+1} = exp(z) g_j
      const double z = rng.normal();       // block 2
      j = jb; np = n + 1; snew = x;
      gnew = exp(z) * g[j];
      const double lu_rev = T->lpr[np][3] + T->logdinv[np];                          // v_0_169 + v_0_219
      const double lu_fwd = (T->lpr[n][2] + log_inv_t) + (-kHalfLog2Pi - 0.5 * z * z);  // v_0_225 + v_0_256 + v_0_263
      lq = (lu_rev - lu_fwd) + log(gnew);                                            // + v_0_446
    } else {                               // coalMoveDeath': delete s_j and g_{j+1}
      np = n - 1;
      const double gd = g[j + 1];
      const double zr = log(gd / g[j]);                                              // v_0_210, v_0_211
      const double lu_rev = (T->lpr[np][2] + log_inv_t) + (-kHalfLog2Pi - 0.5 * zr * zr);  // v_0_170 + v_0_205 + v_0_212
      const double lu_fwd = T->lpr[n][3] + T->logdinv[n];                            // v_0_224 + v_0_266
      lq = (lu_rev - lu_fwd) - log(gd);
    }
    v.set(s, g, move, j, snew, gnew);
    const double fy = eval_f<G>(v, np, ch.ysm, T, k, l, ch.warp_rounds(np));
    const double logr = (fy - lp) + lq;                                              // v_0_448
    const bool accept = uB < fmin(1.0, exp(logr));                                   // x_show_0_8 <- bernoulli (min 1 ratio)
    // commit.  Rate/position moves touch one element; an accepted birth/death shifts the tail: every lane
    // re-reads its elements through the view, the group syncs, then writes in place.
    const bool shift = accept && move >= 3;
    if (__any_sync(0xffffffffu, shift)) {
      if (!shift) v.set(s, g, 0, 0, 0.0, 0.0);
      const int nn = shift ? np : n;
      double vs[ROW / G], vg[ROW / G];
#pragma unroll
      for (int r = 0; r < ROW / G; ++r) {
        const int ic = min(r * G + l + 1, nn + 1);
        vs[r] = v.ps(ic); vg[r] = v.pg(ic);
      }
      __syncwarp();
#pragma unroll
      for (int r = 0; r < ROW / G; ++r) {
        const int i = r * G + l + 1;
        if (i <= nn + 1) { s[i] = vs[r]; g[i] = vg[r]; }
      }
    }
    if (accept && l == 0) {
      if (move == 1) g[j] = gnew;
      if (move == 2) s[j] = snew;
    }
    __syncwarp();
    if (accept) { n = np; lp = fy; ++n_acc; }
    long long slot;
    if (ch.active && record_slot(a, tr, &slot)) {
      if (l == 0) record_value(a, slot, 0, ch.chain, (double)n);
      for (int i = l + 1; i <= SMAX + 1; i += G) {
        if (i <= SMAX) record_value(a, slot, i, ch.chain, i <= n ? s[i] : 0.0);
        record_value(a, slot, SMAX + i, ch.chain, i <= n + 1 ? g[i] : 0.0);
      }
    }
  }
  ch.store(a, n, lp);
  if (ch.active && l == 0 && a.accept) a.accept[ch.chain] += n_acc;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
// v2: `coal_rj1` — ONE lane per chain, built for the LATENCY of a transition (profiles/r01_coal_rj.md: with 4096
// chains the kernel is bound by the dependent chain of one transition, throughput = chains / latency).
//   * per-segment caches in shared memory, [row][lane]: s (upper boundary), F = s/t, cnt = #events below s,
//     g, lg = log g, lw = log(F_i - F_{i-1}).  The joint density is head[n] + sum_i term_i with
//       term_i = lw_i + (alb + am1 lg_i - b g_i) + ((cnt_i - cnt_{i-1}) lg_i - (s_i - s_{i-1}) g_i),
//     and a move touches at most two segments: f(y) - f(x) is evaluated from the <= 2 new and <= 2 old terms (the
//     reference re-evaluates all n+1 segments and 191 events twice per transition).  lp is re-summed in full when the
//     state is stored, so the carried value cannot drift across launches.
//   * one branch-free instruction stream for the four move types (32 different chains share a warp);
//   * log g of a proposed rate is known in closed form (u + log g_j, z + log g_j), so the proposal needs ONE exp,
//     at most two logs (segment widths) and one event-count lookup (bucket table + short scan), all independent;
//   * accept test as log(uB) < log r — the final exp leaves the critical path — with log(uB), the Box-Muller normal
//     and all Philox blocks of transition t+1 computed while transition t waits on its exp/log chains.
// Same Philox layout and the same draws -> proposals as v1 and the oracle; results differ from the full
// re-evaluation only by rounding (~1e-15 relative on log r).
// ---------------------------------------------------------------------------------------------------------
namespace {

constexpr int NB = 256;               // buckets of the event-count table
constexpr int R1 = 34;                // rows per chain: 0 .. 33
constexpr int V2_SMEM_DOUBLES = 196 + TAB_DOUBLES + 5 * R1 * 32;   // y(+pad) | Tabs | s F g lg lw
constexpr int V2_SMEM_BYTES = V2_SMEM_DOUBLES * 8 + R1 * 32 * 4 + (NB + 1) * 4;

struct Draws { double uA, luB, uC, uD, z; };

__device__ __forceinline__ Draws draw5(u64 seed, u64 gchain, u64 step) {
  Rng rng; rng.init(seed, gchain, step);
  Draws d; double uB, u1, u2;
  rng.u01x2(d.uA, uB);
  rng.u01x2(d.uC, d.uD);
  rng.u01x2(u1, u2);                                  // block 2: Box-Muller, cosine branch (Rng::normal)
  d.luB = log(uB);
  double sn, cs; sincospi(2.0 * u2, &sn, &cs);
  d.z = sqrt(-2.0 * log(u1)) * cs;
  return d;
}

// #events strictly below v: bucket table start, then a short scan (y padded with +inf)
__device__ __forceinline__ int count_below_tab(const double* __restrict__ y, const int* __restrict__ tab, int ny,
                                               double v, double nb_over_t) {
  int b = (int)(v * nb_over_t);
  b = max(0, min(b, NB - 1));
  int c = tab[b];
  while (c > 0 && !(y[c - 1] < v)) --c;               // rounding of the bucket index (practically never taken)
  int k = (int)(y[c] < v) + (int)(y[c + 1] < v) + (int)(y[c + 2] < v) + (int)(y[c + 3] < v);
  c += k;
  if (k == 4) while (c < ny && y[c] < v) ++c;
  return c;
}

__device__ __forceinline__ double seg_term(double lw, double lg, double g, int dc, double ds, const Consts& k) {
  return lw + ((k.alb + k.am1 * lg) - k.b * g) + ((double)dc * lg - ds * g);
}

}  // namespace

extern "C" __global__ void __launch_bounds__(32) coal_rj1(const __grid_constant__ StkArgs a) {
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x;
  const int ny = (int)a.flen[F_Y];
  double* ysm = smem;                                   // [ny + 4]
  Tabs* T = reinterpret_cast<Tabs*>(smem + 196);
  double* S = smem + 196 + TAB_DOUBLES;                 // [R1][32] each
  double* Fc = S + R1 * 32; double* Gv = Fc + R1 * 32; double* LG = Gv + R1 * 32; double* LW = LG + R1 * 32;
  int* CN = reinterpret_cast<int*>(LW + R1 * 32);       // [R1][32]
  int* btab = CN + R1 * 32;                             // [NB + 1]
  Consts k;
  k.t = a.par[PAR_T]; k.a = a.par[PAR_A]; k.b = a.par[PAR_B]; k.ny = ny;
  k.alb = k.a * log(k.b) - lgamma(k.a); k.am1 = k.a - 1.0;
  const int lam = (int)a.par[PAR_LAM], nmax = min((int)a.par[PAR_NMAX], SMAX);
  const double* Y = (const double*)a.field[F_Y];
  for (int i = lane; i < ny + 4; i += 32) ysm[i] = i < ny ? Y[i] : pos_inf();
  build_tabs(T, k.t, lam, nmax);
  __syncwarp();
  const double nb_over_t = (double)NB / k.t;
  for (int b = lane; b <= NB; b += 32) btab[b] = count_below(ysm, ny, (double)b * (k.t / (double)NB));
  __syncwarp();

  const long long chain = (long long)blockIdx.x * 32 + lane;
  const bool active = chain < a.n_chains;
  const long long cidx = active ? chain : a.n_chains - 1;
#define AT(A, i) A[(i) * 32 + lane]
  int n = ((const int*)a.field[F_N])[cidx];
  n = max(0, min(n, SMAX));
  {
    const double* Sg = (const double*)a.field[F_S];
    const double* Gg = (const double*)a.field[F_G];
    AT(S, 0) = 0.0; AT(Fc, 0) = 0.0; AT(CN, 0) = 0; AT(Gv, 0) = 0.0; AT(LG, 0) = 0.0; AT(LW, 0) = 0.0;
    for (int i = 1; i < R1; ++i) {
      const double si = i <= n ? Sg[(long long)(i - 1) * a.n_chains + cidx] : k.t;
      const double gi = i <= n + 1 ? Gg[(long long)(i - 1) * a.n_chains + cidx] : 1.0;
      const double Fi = si < 0.0 ? 0.0 : (si > k.t ? 1.0 : si / k.t);
      AT(S, i) = si; AT(Fc, i) = Fi; AT(Gv, i) = gi; AT(LG, i) = log(gi);
      AT(CN, i) = i <= n ? count_below(ysm, ny, si) : ny;
      AT(LW, i) = log(Fi - AT(Fc, i - 1));
    }
  }
  // full evaluation (same left-to-right order as the oracle's fold); also used when the state is stored
  auto full_lp = [&](int nn) {
    double acc = 0.0; bool bad = false;
    for (int i = 1; i <= nn + 1; ++i) {
      acc += seg_term(AT(LW, i), AT(LG, i), AT(Gv, i), AT(CN, i) - AT(CN, i - 1), AT(S, i) - AT(S, i - 1), k);
      if (i <= nn && (AT(S, i) < 0.0 || k.t < AT(S, i))) bad = true;
    }
    return bad ? neg_inf() : acc + T->head[nn];
  };
  double lp = full_lp(n);
  const double log_inv_t = log(1.0 / (k.t - 0.0));
  const u64 gchain = (u64)(a.chain0 + cidx);
  unsigned long long n_acc = 0;
  Draws nx = draw5(a.seed, gchain, a.step0);

  for (long long tr = 0; tr < a.n_steps; ++tr) {
    const Draws d = nx;
    nx = draw5(a.seed, gchain, a.step0 + (u64)tr + 1);   // independent of everything below: fills the stall slots
    // ---- move type (pmf draw), index, birth position -------------------------------------------------
    const double* pr = T->pr[n];
    const double p0 = pr[0], p1 = pr[1], p2 = pr[2], p3 = pr[3];
    const double ua = d.uA * (((p0 + p1) + p2) + p3);
    const double c1 = p0, c2 = c1 + p1, c3 = c2 + p2;
    const int move = ua < c1 ? 1 : (ua < c2 ? 2 : (ua < c3 ? 3 : 4));
    const int m = max((move == 1) ? n + 1 : n, 1);
    int j = (int)floor(d.uC * (double)m); if (j >= m) j = m - 1; j += 1;
    const double x = 0.0 + (k.t - 0.0) * d.uD;
    if (move == 3) {                                     // findSortedInsertIndex x s: first rank with x < s_i
      int lo = 1, hi = n + 1;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (x < AT(S, mid)) hi = mid; else lo = mid + 1; }
      j = lo;
    }
    // ---- neighbourhood of rank j -----------------------------------------------------------------------
    const double sm1 = AT(S, j - 1), s0 = AT(S, j), sp1 = AT(S, j + 1);
    const double Fm1 = AT(Fc, j - 1), F0 = AT(Fc, j), Fp1 = AT(Fc, j + 1);
    const int cm1 = AT(CN, j - 1), c0 = AT(CN, j), cp1 = AT(CN, j + 1);
    const double g0 = AT(Gv, j), lg0 = AT(LG, j), lw0 = AT(LW, j);
    const double g1 = AT(Gv, j + 1), lg1 = AT(LG, j + 1), lw1 = AT(LW, j + 1);
    // ---- the proposal: one exp, one count, two logs, all independent -----------------------------------
    const double u = -0.5 + (0.5 - (-0.5)) * d.uD;                        // coalMoveH
    const double lgn = (move == 1 ? u : d.z) + lg0;                       // log of the proposed rate, closed form
    const double E = exp(move == 1 ? lgn : d.z);
    const double gn = move == 1 ? E : E * g0;                             // exp(u + log g_j)  |  exp(z) g_j
    const double snew = move == 2 ? sm1 + (sp1 - sm1) * (0.0 + (1.0 - 0.0) * d.uD) : x;   // coalMoveP | birth point
    const int cn = count_below_tab(ysm, btab, ny, snew, nb_over_t);
    const double Fn = snew < 0.0 ? 0.0 : (snew > k.t ? 1.0 : snew / k.t);
    const double hiA = move == 4 ? Fp1 : Fn;                              // segment A = (s_{j-1}, snew | s_{j+1})
    const double hiB = move == 2 ? Fp1 : F0;                              // segment B = (snew, s_{j+1} | s_j)
    const double wA = log(hiA - Fm1);
    const double wB = log(hiB - Fn);
    // ---- f(y) - f(x) from the touched segments ------------------------------------------------------------
    const double old0 = seg_term(lw0, lg0, g0, c0 - cm1, s0 - sm1, k);
    const double old1 = seg_term(lw1, lg1, g1, cp1 - c0, sp1 - s0, k);
    double df, lq; int np = n;
    if (move == 1) {
      df = seg_term(lw0, lgn, gn, c0 - cm1, s0 - sm1, k) - old0;
      lq = u;                                                             // log(g_j'/g_j)
    } else if (move == 2) {
      df = (seg_term(wA, lg0, g0, cn - cm1, snew - sm1, k) + seg_term(wB, lg1, g1, cp1 - cn, sp1 - snew, k)) - (old0 + old1);
      lq = 0.0;
    } else if (move == 3) {
      np = n + 1;
      df = ((seg_term(wA, lg0, g0, cn - cm1, snew - sm1, k) + seg_term(wB, lgn, gn, c0 - cn, s0 - snew, k)) - old0)
           + (T->head[np] - T->head[n]);
      const double lu_rev = T->lpr[np][3] + T->logdinv[np];
      const double lu_fwd = (T->lpr[n][2] + log_inv_t) + (-kHalfLog2Pi - 0.5 * d.z * d.z);
      lq = (lu_rev - lu_fwd) + lgn;                                       // + log g'_{j+1}
    } else {
      np = n - 1;
      df = (seg_term(wA, lg0, g0, cp1 - cm1, sp1 - sm1, k) - (old0 + old1)) + (T->head[np] - T->head[n]);
      const double zr = lg1 - lg0;                                        // log(g_{j+1}/g_j)
      const double lu_rev = (T->lpr[np][2] + log_inv_t) + (-kHalfLog2Pi - 0.5 * zr * zr);
      const double lu_fwd = T->lpr[n][3] + T->logdinv[n];
      lq = (lu_rev - lu_fwd) - lg1;                                       // - log g_{j+1}
    }
    const double logr = df + lq;
    const bool accept = d.luB < logr;                                     // uB < min(1, exp(log r))
    // ---- commit ------------------------------------------------------------------------------------------
    if (accept) {
      if (move == 1) { AT(Gv, j) = gn; AT(LG, j) = lgn; }
      else if (move == 2) { AT(S, j) = snew; AT(Fc, j) = Fn; AT(CN, j) = cn; AT(LW, j) = wA; AT(LW, j + 1) = wB; }
      else if (move == 3) {
        for (int i = n + 1; i >= j; --i) { AT(S, i + 1) = AT(S, i); AT(Fc, i + 1) = AT(Fc, i); AT(CN, i + 1) = AT(CN, i); }
        for (int i = n + 1; i > j; --i) { AT(Gv, i + 1) = AT(Gv, i); AT(LG, i + 1) = AT(LG, i); AT(LW, i + 1) = AT(LW, i); }
        AT(S, j) = snew; AT(Fc, j) = Fn; AT(CN, j) = cn; AT(LW, j) = wA;
        AT(Gv, j + 1) = gn; AT(LG, j + 1) = lgn; AT(LW, j + 1) = wB;
      } else {
        for (int i = j; i <= n; ++i) { AT(S, i) = AT(S, i + 1); AT(Fc, i) = AT(Fc, i + 1); AT(CN, i) = AT(CN, i + 1); }
        for (int i = j + 1; i <= n; ++i) { AT(Gv, i) = AT(Gv, i + 1); AT(LG, i) = AT(LG, i + 1); AT(LW, i) = AT(LW, i + 1); }
        AT(LW, j) = wA;
      }
      n = np; lp += df; ++n_acc;
    }
    long long slot;
    if (active && record_slot(a, tr, &slot)) {
      record_value(a, slot, 0, chain, (double)n);
      for (int i = 1; i <= SMAX + 1; ++i) {
        if (i <= SMAX) record_value(a, slot, i, chain, i <= n ? AT(S, i) : 0.0);
        record_value(a, slot, SMAX + i, chain, i <= n + 1 ? AT(Gv, i) : 0.0);
      }
    }
  }
  if (active) {
    double* Sg = (double*)a.field[F_S];
    double* Gg = (double*)a.field[F_G];
    for (int i = 1; i <= SMAX + 1; ++i) {
      if (i <= SMAX) Sg[(long long)(i - 1) * a.n_chains + chain] = i <= n ? AT(S, i) : 0.0;
      Gg[(long long)(i - 1) * a.n_chains + chain] = i <= n + 1 ? AT(Gv, i) : 0.0;
    }
    ((int*)a.field[F_N])[chain] = n;
    ((double*)a.field[F_LP])[chain] = full_lp(n);
    if (a.accept) a.accept[chain] += n_acc;
  }
#undef AT
}
static_assert(V2_SMEM_BYTES == 53044, "V2_SMEM_BYTES");

// ---------------------------------------------------------------------------------------------------------
// v3: `coal_draws` + `coal_rj2<LPW>` — the state-independent half of a transition (3 Philox blocks, log uB, the
// Box-Muller normal: ~600 of v2's ~1150 instructions) moves into a PRE-PASS kernel that runs one thread per
// (chain, transition) at full occupancy and leaves [transition][5][chain] doubles in HBM; the sequential kernel
// streams them back (coalesced, prefetched two transitions ahead) and keeps only the dependent chain:
// select move -> neighbourhood (packed 16-byte rows) -> one exp | one count | two logs -> compare -> commit.
// LPW = chains per warp: with 4096 chains the kernel is latency-bound whatever the packing, and fewer chains per
// warp means fewer lanes to wait for in the divergent parts (rank search of a birth, row shifts of an accepted
// birth/death).  Same draws -> proposals as v1/v2 and the oracle.
// ---------------------------------------------------------------------------------------------------------
extern "C" __global__ void __launch_bounds__(256) coal_draws(const __grid_constant__ StkArgs a) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.n_chains * a.n_steps) return;
  const long long tr = idx / a.n_chains, c = idx - tr * a.n_chains;
  const Draws d = draw5(a.seed, (u64)(a.chain0 + c), a.step0 + (u64)tr);
  // [transition][chain][8]: everything about the transition that does not depend on the chain's state, including
  // the two exps of the rate proposals (g' = exp(u) g_j for the rate move, exp(z) g_j for a birth)
  double2* o = reinterpret_cast<double2*>(a.scratch) + (tr * a.n_chains + c) * 4;
  const double u = -0.5 + (0.5 - (-0.5)) * d.uD;
  o[0] = make_double2(d.uA, d.luB);
  o[1] = make_double2(d.uC, d.uD);
  o[2] = make_double2(d.z, 0.5 * d.z * d.z);
  o[3] = make_double2(exp(u), exp(d.z));
}

namespace {

struct Tabs2 { double birthC[32]; double deathC[32]; };   // per-n constants of log q ratios + head differences

template <int LPW>
__device__ __forceinline__ void rj2_body(const StkArgs& a) {
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x;
  const int ny = (int)a.flen[F_Y];
  double* ysm = smem;                                   // [ny + 4]
  Tabs* T = reinterpret_cast<Tabs*>(smem + 196);
  Tabs2* T2 = reinterpret_cast<Tabs2*>(smem + 196 + TAB_DOUBLES);
  double* fm = smem + 196 + TAB_DOUBLES + 64;           // fast exp/log tables (stk_fastmath.cuh)
  double2* P = reinterpret_cast<double2*>(fm + kFmTableDoubles);             // [R1][LPW] (s, #events below s)
  double2* Q = P + R1 * LPW;                                                 // (g, log g)
  double* LW = reinterpret_cast<double*>(Q + R1 * LPW);                      // log((s_i - s_{i-1}) / t)
  double2* ring = reinterpret_cast<double2*>(LW + R1 * LPW + ((R1 * LPW) & 1));   // [4][LPW][4] double2: draws in flight
  int* btab = reinterpret_cast<int*>(ring + 4 * LPW * 4);                    // [NB + 1]
  Consts k;
  k.t = a.par[PAR_T]; k.a = a.par[PAR_A]; k.b = a.par[PAR_B]; k.ny = ny;
  k.alb = k.a * log(k.b) - lgamma(k.a); k.am1 = k.a - 1.0;
  const int lam = (int)a.par[PAR_LAM], nmax = min((int)a.par[PAR_NMAX], SMAX);
  const double* Y = (const double*)a.field[F_Y];
  for (int i = lane; i < ny + 4; i += 32) ysm[i] = i < ny ? Y[i] : pos_inf();
  build_tabs(T, k.t, lam, nmax);
  fm_build_tables(fm, lane, 32);
  __syncwarp();
  const double log_inv_t = log(1.0 / (k.t - 0.0));
  const double inv_t = 1.0 / k.t;
  {
    const int n = lane;                                  // one n per lane
    const int nu = min(n + 1, 31), nd = max(n - 1, 0);
    T2->birthC[n] = ((T->lpr[nu][3] + T->logdinv[nu]) - ((T->lpr[n][2] + log_inv_t) - kHalfLog2Pi)) + (T->head[nu] - T->head[n]);
    T2->deathC[n] = (((T->lpr[nd][2] + log_inv_t) - kHalfLog2Pi) - (T->lpr[n][3] + T->logdinv[n])) + (T->head[nd] - T->head[n]);
  }
  const double nb_over_t = (double)NB / k.t;
  for (int b = lane; b <= NB; b += 32) btab[b] = count_below(ysm, ny, (double)b * (k.t / (double)NB));
  __syncwarp();
  if (lane >= LPW) return;

  const long long chain = (long long)blockIdx.x * LPW + lane;
  const bool active = chain < a.n_chains;
  const long long cidx = active ? chain : a.n_chains - 1;
#define ROW(A, i) A[(i) * LPW + lane]
  int n = ((const int*)a.field[F_N])[cidx];
  n = max(0, min(n, SMAX));
  {
    const double* Sg = (const double*)a.field[F_S];
    const double* Gg = (const double*)a.field[F_G];
    ROW(P, 0) = make_double2(0.0, 0.0); ROW(Q, 0) = make_double2(0.0, 0.0); ROW(LW, 0) = 0.0;
    double sprev = 0.0;
    for (int i = 1; i < R1; ++i) {
      const double si = i <= n ? Sg[(long long)(i - 1) * a.n_chains + cidx] : k.t;
      const double gi = i <= n + 1 ? Gg[(long long)(i - 1) * a.n_chains + cidx] : 1.0;
      ROW(P, i) = make_double2(si, (double)(i <= n ? count_below(ysm, ny, si) : ny));
      ROW(Q, i) = make_double2(gi, log(gi));
      ROW(LW, i) = log((si - sprev) * inv_t);
      sprev = si;
    }
  }
  auto full_lp = [&](int nn) {
    double acc = 0.0; bool bad = false;
    for (int i = 1; i <= nn + 1; ++i) {
      const double2 p1 = ROW(P, i), p0 = ROW(P, i - 1), q = ROW(Q, i);
      acc += ROW(LW, i) + ((k.alb + k.am1 * q.y) - k.b * q.x) + ((p1.y - p0.y) * q.y - (p1.x - p0.x) * q.x);
      if (i <= nn && (p1.x < 0.0 || k.t < p1.x)) bad = true;
    }
    return bad ? neg_inf() : acc + T->head[nn];
  };
  double lp = full_lp(n);
  unsigned long long n_acc = 0;
  // draws of this chain: [tr][chain][8] in HBM -> a 4-deep cp.async ring in shared memory (three transitions in
  // flight; a register FIFO would be rotated by the compiler right behind the loads and wait for them)
  const double2* D = reinterpret_cast<const double2*>(a.scratch) + cidx * 4;
  const long long NC4 = a.n_chains * 4;
  const unsigned ring_base = (unsigned)__cvta_generic_to_shared(ring + lane * 4);
  auto fetch = [&](long long tr) {
    if (tr < a.n_steps) {
      const double2* p = D + tr * NC4;
      const unsigned dst = ring_base + (unsigned)((tr & 3) * LPW * 64);
#pragma unroll
      for (int q = 0; q < 4; ++q)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + q * 16), "l"(p + q) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  fetch(0); fetch(1); fetch(2);
  const bool rec = active && a.draws && a.draw_slots > 0;

  for (long long tr = 0; tr < a.n_steps; ++tr) {
    fetch(tr + 3);
    asm volatile("cp.async.wait_group 3;" ::: "memory");
    const double2* dr = ring + ((tr & 3) * LPW + lane) * 4;
    const double2 dA = dr[0], dC = dr[1], dZ = dr[2], dE = dr[3];
    const double uA = dA.x, luB = dA.y, uC = dC.x, uD = dC.y, z = dZ.x, hz2 = dZ.y;
    // ---- move type (pmf draw), index, birth position -------------------------------------------------
    const double2 pa = *reinterpret_cast<const double2*>(&T->pr[n][0]), pb = *reinterpret_cast<const double2*>(&T->pr[n][2]);
    const double ua = uA * (((pa.x + pa.y) + pb.x) + pb.y);
    const double c1 = pa.x, c2 = c1 + pa.y, c3 = c2 + pb.x;
    const int move = ua < c1 ? 1 : (ua < c2 ? 2 : (ua < c3 ? 3 : 4));
    const int m = max((move == 1) ? n + 1 : n, 1);
    int j = (int)(uC * (double)m); if (j >= m) j = m - 1; j += 1;
    const double x = 0.0 + (k.t - 0.0) * uD;
    if (move == 3) {                                     // findSortedInsertIndex x s: first rank with x < s_i
      int lo = 1, hi = n + 1;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (x < ROW(P, mid).x) hi = mid; else lo = mid + 1; }
      j = lo;
    }
    // ---- neighbourhood of rank j -----------------------------------------------------------------------
    const double2 Pm1 = ROW(P, j - 1), P0 = ROW(P, j), Pp1 = ROW(P, j + 1);
    const double2 Q0 = ROW(Q, j), Q1 = ROW(Q, j + 1);
    const double lw0 = ROW(LW, j), lw1 = ROW(LW, j + 1);
    const double bC = T2->birthC[n], dCn = T2->deathC[n];
    const double sm1 = Pm1.x, s0 = P0.x, sp1 = Pp1.x, cm1 = Pm1.y, c0 = P0.y, cp1 = Pp1.y;
    const double g0 = Q0.x, lg0 = Q0.y, g1 = Q1.x, lg1 = Q1.y;
    // ---- the proposal: one exp, one count, two logs, all independent -----------------------------------
    const bool m1 = move == 1, m2 = move == 2, m3 = move == 3, m4 = move == 4;
    const double u = -0.5 + (0.5 - (-0.5)) * uD;                          // coalMoveH
    const double lgn = (m1 ? u : z) + lg0;                                // log of the proposed rate, closed form
    const double gn = (m1 ? dE.x : dE.y) * g0;                            // exp(u) g_j  |  exp(z) g_j (exps from the pre-pass)
    const double snew = m2 ? sm1 + (sp1 - sm1) * (0.0 + (1.0 - 0.0) * uD) : x;   // coalMoveP | birth point
    const double cn = (double)count_below_tab(ysm, btab, ny, snew, nb_over_t);
    const double hiA = m1 ? s0 : (m4 ? sp1 : snew), hiB = m2 ? sp1 : s0;
    const double dsA = hiA - sm1, dsB = hiB - snew;                       // widths of segment A = (s_{j-1}, .), B = (snew, .)
    const double wA = fast_log(dsA * inv_t, fm);
    const double wB = fast_log(dsB * inv_t, fm);
    // ---- f(y) - f(x) from the touched segments (selects, no branches) ----------------------------------
    const double old0 = seg_term(lw0, lg0, g0, 0, 0.0, k) + ((c0 - cm1) * lg0 - (s0 - sm1) * g0);
    const double old1 = seg_term(lw1, lg1, g1, 0, 0.0, k) + ((cp1 - c0) * lg1 - (sp1 - s0) * g1);
    const double chA = m1 ? c0 : (m4 ? cp1 : cn);
    const double lgA = m1 ? lgn : lg0, gA = m1 ? gn : g0, lwA = m1 ? lw0 : wA;
    const double tA = seg_term(lwA, lgA, gA, 0, 0.0, k) + ((chA - cm1) * lgA - dsA * gA);
    const double chB = m2 ? cp1 : c0, lgB = m2 ? lg1 : lgn, gB = m2 ? g1 : gn;
    const double tB = (m2 || m3) ? seg_term(wB, lgB, gB, 0, 0.0, k) + ((chB - cn) * lgB - dsB * gB) : 0.0;
    const double told = (m2 || m4) ? old0 + old1 : old0;
    const double zr = lg1 - lg0;                                          // log(g_{j+1}/g_j), death
    const double extra = m1 ? u : (m2 ? 0.0 : (m3 ? (bC + hz2) + lgn : (dCn - 0.5 * zr * zr) - lg1));
    const double df = (tA + tB) - told;                                   // without the head difference (in bC/dC)
    const double logr = df + extra;
    const bool accept = luB < logr;                                       // uB < min(1, exp(log r))
    // ---- commit ------------------------------------------------------------------------------------------
    if (accept) {
      if (m1) ROW(Q, j) = make_double2(gn, lgn);
      else if (m2) { ROW(P, j) = make_double2(snew, cn); ROW(LW, j) = wA; ROW(LW, j + 1) = wB; }
      else if (m3) {
        for (int i = n + 1; i >= j; --i) {               // rows j.. move up; row j is then re-written
          const double2 p = ROW(P, i), q = ROW(Q, i); const double w = ROW(LW, i);
          ROW(P, i + 1) = p; ROW(Q, i + 1) = q; ROW(LW, i + 1) = w;
        }
        ROW(P, j) = make_double2(snew, cn); ROW(LW, j) = wA;
        ROW(Q, j + 1) = make_double2(gn, lgn); ROW(LW, j + 1) = wB;
        ++n;
      } else {
        for (int i = j; i <= n; ++i) {                   // rows j+1.. move down; row j keeps its rate, new width
          const double2 p = ROW(P, i + 1), q = ROW(Q, i + 1); const double w = ROW(LW, i + 1);
          ROW(P, i) = p; ROW(Q, i) = q; ROW(LW, i) = w;
        }
        ROW(Q, j) = make_double2(g0, lg0); ROW(LW, j) = wA;
        --n;
      }
      lp += df + (m3 ? T->head[n] - T->head[n - 1] : (m4 ? T->head[n] - T->head[n + 1] : 0.0));
      ++n_acc;
    }
    long long slot;
    if (rec && record_slot(a, tr, &slot)) {
      record_value(a, slot, 0, chain, (double)n);
      for (int i = 1; i <= SMAX + 1; ++i) {
        if (i <= SMAX) record_value(a, slot, i, chain, i <= n ? ROW(P, i).x : 0.0);
        record_value(a, slot, SMAX + i, chain, i <= n + 1 ? ROW(Q, i).x : 0.0);
      }
    }
  }
  if (active) {
    double* Sg = (double*)a.field[F_S];
    double* Gg = (double*)a.field[F_G];
    for (int i = 1; i <= SMAX + 1; ++i) {
      if (i <= SMAX) Sg[(long long)(i - 1) * a.n_chains + chain] = i <= n ? ROW(P, i).x : 0.0;
      Gg[(long long)(i - 1) * a.n_chains + chain] = i <= n + 1 ? ROW(Q, i).x : 0.0;
    }
    ((int*)a.field[F_N])[chain] = n;
    ((double*)a.field[F_LP])[chain] = full_lp(n);
    if (a.accept) a.accept[chain] += n_acc;
  }
#undef ROW
}

// ---------------------------------------------------------------------------------------------------------
// v4: `coal_rj3<LPW, CPL>` — CPL chains per LANE (round 2).  v3 is bound by the in-order issue of ONE dependent
// instruction stream per warp (≈ 440 instructions at 4.95 cycles each for a true dependent chain of ≈ 500 cycles,
// 1.7 warps per scheduler at 4096 chains): the schedulers have nothing else to issue.  Here every lane carries CPL
// independent chains and a transition of all of them is written as two phases: phase A (select move, neighbourhood
// loads, proposal, the two logs, log r, the decision) for all CPL chains with NO shared-memory store in between, so
// that the compiler interleaves the CPL independent streams (instruction-level parallelism instead of warps); then
// phase B (commit, record).  Same draws, same arithmetic per chain as v3: bit-identical trajectories.
// Chain of a lane: local index c * LPW + lane of the CTA's LPW * CPL chains.
// ---------------------------------------------------------------------------------------------------------
template <int LPW, int CPL>
__device__ __forceinline__ void rj3_body(const StkArgs& a) {
  constexpr int CPW = LPW * CPL;
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x;
  const int ny = (int)a.flen[F_Y];
  double* ysm = smem;                                   // [ny + 4]
  Tabs* T = reinterpret_cast<Tabs*>(smem + 196);
  Tabs2* T2 = reinterpret_cast<Tabs2*>(smem + 196 + TAB_DOUBLES);
  double* fm = smem + 196 + TAB_DOUBLES + 64;
  double2* P = reinterpret_cast<double2*>(fm + kFmTableDoubles);             // [R1][CPW] (s, #events below s)
  double2* Q = P + R1 * CPW;                                                 // (g, log g)
  double* LW = reinterpret_cast<double*>(Q + R1 * CPW);                      // log((s_i - s_{i-1}) / t)
  double2* ring = reinterpret_cast<double2*>(LW + R1 * CPW + ((R1 * CPW) & 1));   // [4][CPW][4] double2: draws in flight
  int* btab = reinterpret_cast<int*>(ring + 4 * CPW * 4);                    // [NB + 1]
  Consts k;
  k.t = a.par[PAR_T]; k.a = a.par[PAR_A]; k.b = a.par[PAR_B]; k.ny = ny;
  k.alb = k.a * log(k.b) - lgamma(k.a); k.am1 = k.a - 1.0;
  const int lam = (int)a.par[PAR_LAM], nmax = min((int)a.par[PAR_NMAX], SMAX);
  const double* Y = (const double*)a.field[F_Y];
  for (int i = lane; i < ny + 4; i += 32) ysm[i] = i < ny ? Y[i] : pos_inf();
  build_tabs(T, k.t, lam, nmax);
  fm_build_tables(fm, lane, 32);
  __syncwarp();
  const double log_inv_t = log(1.0 / (k.t - 0.0));
  const double inv_t = 1.0 / k.t;
  {
    const int n = lane;
    const int nu = min(n + 1, 31), nd = max(n - 1, 0);
    T2->birthC[n] = ((T->lpr[nu][3] + T->logdinv[nu]) - ((T->lpr[n][2] + log_inv_t) - kHalfLog2Pi)) + (T->head[nu] - T->head[n]);
    T2->deathC[n] = (((T->lpr[nd][2] + log_inv_t) - kHalfLog2Pi) - (T->lpr[n][3] + T->logdinv[n])) + (T->head[nd] - T->head[n]);
  }
  const double nb_over_t = (double)NB / k.t;
  for (int b = lane; b <= NB; b += 32) btab[b] = count_below(ysm, ny, (double)b * (k.t / (double)NB));
  __syncwarp();
  if (lane >= LPW) return;

#define ROWC(A, i, c) A[(i) * CPW + (c) * LPW + lane]
  long long chain[CPL], cidx[CPL]; bool active[CPL]; int n[CPL]; double lp[CPL]; unsigned long long n_acc[CPL];
  auto full_lp = [&](int c, int nn) {
    double acc = 0.0; bool bad = false;
    for (int i = 1; i <= nn + 1; ++i) {
      const double2 p1 = ROWC(P, i, c), p0 = ROWC(P, i - 1, c), q = ROWC(Q, i, c);
      acc += ROWC(LW, i, c) + ((k.alb + k.am1 * q.y) - k.b * q.x) + ((p1.y - p0.y) * q.y - (p1.x - p0.x) * q.x);
      if (i <= nn && (p1.x < 0.0 || k.t < p1.x)) bad = true;
    }
    return bad ? neg_inf() : acc + T->head[nn];
  };
#pragma unroll
  for (int c = 0; c < CPL; ++c) {
    chain[c] = (long long)blockIdx.x * CPW + c * LPW + lane;
    active[c] = chain[c] < a.n_chains;
    cidx[c] = active[c] ? chain[c] : a.n_chains - 1;
    int nn = ((const int*)a.field[F_N])[cidx[c]];
    nn = max(0, min(nn, SMAX));
    n[c] = nn;
    const double* Sg = (const double*)a.field[F_S];
    const double* Gg = (const double*)a.field[F_G];
    ROWC(P, 0, c) = make_double2(0.0, 0.0); ROWC(Q, 0, c) = make_double2(0.0, 0.0); ROWC(LW, 0, c) = 0.0;
    double sprev = 0.0;
    for (int i = 1; i < R1; ++i) {
      const double si = i <= nn ? Sg[(long long)(i - 1) * a.n_chains + cidx[c]] : k.t;
      const double gi = i <= nn + 1 ? Gg[(long long)(i - 1) * a.n_chains + cidx[c]] : 1.0;
      ROWC(P, i, c) = make_double2(si, (double)(i <= nn ? count_below(ysm, ny, si) : ny));
      ROWC(Q, i, c) = make_double2(gi, log(gi));
      ROWC(LW, i, c) = log((si - sprev) * inv_t);
      sprev = si;
    }
    lp[c] = full_lp(c, nn);
    n_acc[c] = 0;
  }
  const long long NC4 = a.n_chains * 4;
  auto fetch = [&](long long tr) {
    if (tr < a.n_steps) {
#pragma unroll
      for (int c = 0; c < CPL; ++c) {
        const double2* p = reinterpret_cast<const double2*>(a.scratch) + cidx[c] * 4 + tr * NC4;
        const unsigned dst = (unsigned)__cvta_generic_to_shared(ring + (((tr & 3) * CPW) + c * LPW + lane) * 4);
#pragma unroll
        for (int q = 0; q < 4; ++q)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + q * 16), "l"(p + q) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  fetch(0); fetch(1); fetch(2);
  bool rec[CPL];
#pragma unroll
  for (int c = 0; c < CPL; ++c) rec[c] = active[c] && a.draws && a.draw_slots > 0;

  for (long long tr = 0; tr < a.n_steps; ++tr) {
    fetch(tr + 3);
    asm volatile("cp.async.wait_group 3;" ::: "memory");
    // ---- phase A: everything up to the decision, for all CPL chains, without a shared-memory store.  Written as a sequence of
    // short sub-phases, each looping over the CPL chains: ptxas schedules inside basic blocks, so the rare-path branches (rank
    // search of a birth, event-count scan, fast_log's fallback) are kept in sub-phases of their own and the long fp64 chains
    // (two logs, four segment terms, log r) of all the chains sit side by side in branch-free blocks. ------------------------
    int mv[CPL], jj[CPL]; bool acc[CPL];
    double gnv[CPL], lgnv[CPL], snewv[CPL], cnv[CPL], wAv[CPL], wBv[CPL], g0v[CPL], lg0v[CPL], dfv[CPL];
    double luBv[CPL], uDv[CPL], zv[CPL], hz2v[CPL], xv[CPL], eAv[CPL], eBv[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) {                      // A1: draws, move type, index
      const double2* dr = ring + (((tr & 3) * CPW) + c * LPW + lane) * 4;
      const double2 dA = dr[0], dC = dr[1], dZ = dr[2], dE = dr[3];
      const int nn = n[c];
      const double2 pa = *reinterpret_cast<const double2*>(&T->pr[nn][0]), pb = *reinterpret_cast<const double2*>(&T->pr[nn][2]);
      const double ua = dA.x * (((pa.x + pa.y) + pb.x) + pb.y);
      const double c1 = pa.x, c2 = c1 + pa.y, c3 = c2 + pb.x;
      const int move = ua < c1 ? 1 : (ua < c2 ? 2 : (ua < c3 ? 3 : 4));
      const int m = max((move == 1) ? nn + 1 : nn, 1);
      int j = (int)(dC.x * (double)m); if (j >= m) j = m - 1; j += 1;
      mv[c] = move; jj[c] = j; luBv[c] = dA.y; uDv[c] = dC.y; zv[c] = dZ.x; hz2v[c] = dZ.y; eAv[c] = dE.x; eBv[c] = dE.y;
      xv[c] = 0.0 + (k.t - 0.0) * dC.y;
    }
#pragma unroll
    for (int c = 0; c < CPL; ++c) if (mv[c] == 3) {      // A2: findSortedInsertIndex x s (birth only)
      int lo = 1, hi = n[c] + 1;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (xv[c] < ROWC(P, mid, c).x) hi = mid; else lo = mid + 1; }
      jj[c] = lo;
    }
    double sm1v[CPL], s0v[CPL], sp1v[CPL], cm1v[CPL], c0v[CPL], cp1v[CPL], g1v[CPL], lg1v[CPL], lw0v[CPL], lw1v[CPL], dsAv[CPL], dsBv[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) {                      // A3: neighbourhood of rank j, the proposal
      const int j = jj[c];
      const double2 Pm1 = ROWC(P, j - 1, c), P0 = ROWC(P, j, c), Pp1 = ROWC(P, j + 1, c);
      const double2 Q0 = ROWC(Q, j, c), Q1 = ROWC(Q, j + 1, c);
      lw0v[c] = ROWC(LW, j, c); lw1v[c] = ROWC(LW, j + 1, c);
      sm1v[c] = Pm1.x; s0v[c] = P0.x; sp1v[c] = Pp1.x; cm1v[c] = Pm1.y; c0v[c] = P0.y; cp1v[c] = Pp1.y;
      g0v[c] = Q0.x; lg0v[c] = Q0.y; g1v[c] = Q1.x; lg1v[c] = Q1.y;
      const bool m1 = mv[c] == 1, m2 = mv[c] == 2, m4 = mv[c] == 4;
      const double u = -0.5 + (0.5 - (-0.5)) * uDv[c];
      lgnv[c] = (m1 ? u : zv[c]) + lg0v[c];
      gnv[c] = (m1 ? eAv[c] : eBv[c]) * g0v[c];
      snewv[c] = m2 ? sm1v[c] + (sp1v[c] - sm1v[c]) * (0.0 + (1.0 - 0.0) * uDv[c]) : xv[c];
      const double hiA = m1 ? s0v[c] : (m4 ? sp1v[c] : snewv[c]), hiB = m2 ? sp1v[c] : s0v[c];
      dsAv[c] = hiA - sm1v[c]; dsBv[c] = hiB - snewv[c];
    }
    {                                                    // A4: events below the new point: bucket start + four probes, branch-free;
      int cb[CPL]; bool rare = false;                    // the rare paths (bucket rounding, > 4 events in the scan) fall back to the loop version
#pragma unroll
      for (int c = 0; c < CPL; ++c) { int b = (int)(snewv[c] * nb_over_t); b = max(0, min(b, NB - 1)); cb[c] = btab[b]; }
#pragma unroll
      for (int c = 0; c < CPL; ++c) {
        const double v = snewv[c]; const int c0i = cb[c];
        rare = rare || (c0i > 0 && !(ysm[c0i - 1] < v));
        const int kq = (int)(ysm[c0i] < v) + (int)(ysm[c0i + 1] < v) + (int)(ysm[c0i + 2] < v) + (int)(ysm[c0i + 3] < v);
        rare = rare || kq == 4;
        cnv[c] = (double)(c0i + kq);
      }
      if (rare) {
#pragma unroll
        for (int c = 0; c < CPL; ++c) cnv[c] = (double)count_below_tab(ysm, btab, ny, snewv[c], nb_over_t);
      }
    }
    {                                                    // A5: the two logs of every chain, one branch-free block
      double la[CPL], lb[CPL]; bool oka = true;
#pragma unroll
      for (int c = 0; c < CPL; ++c) { la[c] = dsAv[c] * inv_t; lb[c] = dsBv[c] * inv_t; oka = oka && fast_log_ok(la[c]) && fast_log_ok(lb[c]); }
#pragma unroll
      for (int c = 0; c < CPL; ++c) { wAv[c] = fast_log_core(la[c], fm); wBv[c] = fast_log_core(lb[c], fm); }
      if (!oka) {                                        // a zero / negative / non-finite width (an impossible proposal): the library's answer
#pragma unroll
        for (int c = 0; c < CPL; ++c) { wAv[c] = fast_log(la[c], fm); wBv[c] = fast_log(lb[c], fm); }
      }
    }
#pragma unroll
    for (int c = 0; c < CPL; ++c) {                      // A6: f(y) - f(x) from the touched segments, log r, the decision
      const bool m1 = mv[c] == 1, m2 = mv[c] == 2, m3 = mv[c] == 3, m4 = mv[c] == 4;
      const double sm1 = sm1v[c], s0 = s0v[c], sp1 = sp1v[c], cm1 = cm1v[c], c0 = c0v[c], cp1 = cp1v[c];
      const double g0 = g0v[c], lg0 = lg0v[c], g1 = g1v[c], lg1 = lg1v[c], lgn = lgnv[c], gn = gnv[c], cn = cnv[c], wA = wAv[c], wB = wBv[c];
      const double u = -0.5 + (0.5 - (-0.5)) * uDv[c];
      const double bC = T2->birthC[n[c]], dCn = T2->deathC[n[c]];
      const double old0 = seg_term(lw0v[c], lg0, g0, 0, 0.0, k) + ((c0 - cm1) * lg0 - (s0 - sm1) * g0);
      const double old1 = seg_term(lw1v[c], lg1, g1, 0, 0.0, k) + ((cp1 - c0) * lg1 - (sp1 - s0) * g1);
      const double chA = m1 ? c0 : (m4 ? cp1 : cn);
      const double lgA = m1 ? lgn : lg0, gA = m1 ? gn : g0, lwA = m1 ? lw0v[c] : wA;
      const double tA = seg_term(lwA, lgA, gA, 0, 0.0, k) + ((chA - cm1) * lgA - dsAv[c] * gA);
      const double chB = m2 ? cp1 : c0, lgB = m2 ? lg1 : lgn, gB = m2 ? g1 : gn;
      const double tB = (m2 || m3) ? seg_term(wB, lgB, gB, 0, 0.0, k) + ((chB - cn) * lgB - dsBv[c] * gB) : 0.0;
      const double told = (m2 || m4) ? old0 + old1 : old0;
      const double zr = lg1 - lg0;
      const double extra = m1 ? u : (m2 ? 0.0 : (m3 ? (bC + hz2v[c]) + lgn : (dCn - 0.5 * zr * zr) - lg1));
      const double df = (tA + tB) - told;
      const double logr = df + extra;
      acc[c] = luBv[c] < logr;
      dfv[c] = df;
    }
    // ---- phase B: commit + record ---------------------------------------------------------------------------------
#pragma unroll
    for (int c = 0; c < CPL; ++c) {
      const int j = jj[c];
      if (acc[c]) {
        const int move = mv[c];
        int nn = n[c];
        if (move == 1) ROWC(Q, j, c) = make_double2(gnv[c], lgnv[c]);
        else if (move == 2) { ROWC(P, j, c) = make_double2(snewv[c], cnv[c]); ROWC(LW, j, c) = wAv[c]; ROWC(LW, j + 1, c) = wBv[c]; }
        else if (move == 3) {
          for (int i = nn + 1; i >= j; --i) {
            const double2 p = ROWC(P, i, c), q = ROWC(Q, i, c); const double w = ROWC(LW, i, c);
            ROWC(P, i + 1, c) = p; ROWC(Q, i + 1, c) = q; ROWC(LW, i + 1, c) = w;
          }
          ROWC(P, j, c) = make_double2(snewv[c], cnv[c]); ROWC(LW, j, c) = wAv[c];
          ROWC(Q, j + 1, c) = make_double2(gnv[c], lgnv[c]); ROWC(LW, j + 1, c) = wBv[c];
          ++nn;
        } else {
          for (int i = j; i <= nn; ++i) {
            const double2 p = ROWC(P, i + 1, c), q = ROWC(Q, i + 1, c); const double w = ROWC(LW, i + 1, c);
            ROWC(P, i, c) = p; ROWC(Q, i, c) = q; ROWC(LW, i, c) = w;
          }
          ROWC(Q, j, c) = make_double2(g0v[c], lg0v[c]); ROWC(LW, j, c) = wAv[c];
          --nn;
        }
        n[c] = nn;
        lp[c] += dfv[c] + (move == 3 ? T->head[nn] - T->head[nn - 1] : (move == 4 ? T->head[nn] - T->head[nn + 1] : 0.0));
        ++n_acc[c];
      }
      long long slot;
      if (rec[c] && record_slot(a, tr, &slot)) {
        record_value(a, slot, 0, chain[c], (double)n[c]);
        for (int i = 1; i <= SMAX + 1; ++i) {
          if (i <= SMAX) record_value(a, slot, i, chain[c], i <= n[c] ? ROWC(P, i, c).x : 0.0);
          record_value(a, slot, SMAX + i, chain[c], i <= n[c] + 1 ? ROWC(Q, i, c).x : 0.0);
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < CPL; ++c) if (active[c]) {
    double* Sg = (double*)a.field[F_S];
    double* Gg = (double*)a.field[F_G];
    for (int i = 1; i <= SMAX + 1; ++i) {
      if (i <= SMAX) Sg[(long long)(i - 1) * a.n_chains + chain[c]] = i <= n[c] ? ROWC(P, i, c).x : 0.0;
      Gg[(long long)(i - 1) * a.n_chains + chain[c]] = i <= n[c] + 1 ? ROWC(Q, i, c).x : 0.0;
    }
    ((int*)a.field[F_N])[chain[c]] = n[c];
    ((double*)a.field[F_LP])[chain[c]] = full_lp(c, n[c]);
    if (a.accept) a.accept[chain[c]] += n_acc[c];
  }
#undef ROWC
}

// smem: (196 + TAB_DOUBLES + 64 + fast-math tables) doubles + rows (40 B per row and chain) + ring + (NB+1) ints
constexpr int rj2_smem(int lpw) { return (196 + TAB_DOUBLES + 64 + kFmTableDoubles) * 8 + R1 * lpw * 40 + 8 + 4 * lpw * 64 + (NB + 1) * 4; }
static_assert(rj2_smem(8) == 23740 && rj2_smem(4) == 17276 && rj2_smem(16) == 36668 && rj2_smem(32) == 62524, "rj2 smem");

}  // namespace

extern "C" __global__ void __launch_bounds__(32) coal_rj2_8(const __grid_constant__ StkArgs a) { rj2_body<8>(a); }
extern "C" __global__ void __launch_bounds__(32) coal_rj2_4(const __grid_constant__ StkArgs a) { rj2_body<4>(a); }
extern "C" __global__ void __launch_bounds__(32) coal_rj2_2(const __grid_constant__ StkArgs a) { rj2_body<2>(a); }
extern "C" __global__ void __launch_bounds__(32) coal_rj2_1(const __grid_constant__ StkArgs a) { rj2_body<1>(a); }
extern "C" __global__ void __launch_bounds__(32) coal_rj2_16(const __grid_constant__ StkArgs a) { rj2_body<16>(a); }
extern "C" __global__ void __launch_bounds__(32) coal_rj2_32(const __grid_constant__ StkArgs a) { rj2_body<32>(a); }
// v4: <lanes per warp> x <chains per lane>; shared memory = rj2_smem(lanes * chains per lane)
extern "C" __global__ void __launch_bounds__(32) coal_rj3_4x2(const __grid_constant__ StkArgs a) { rj3_body<4, 2>(a); }
extern "C" __global__ void __launch_bounds__(32) coal_rj3_2x2(const __grid_constant__ StkArgs a) { rj3_body<2, 2>(a); }
extern "C" __global__ void __launch_bounds__(32) coal_rj3_2x4(const __grid_constant__ StkArgs a) { rj3_body<2, 4>(a); }
extern "C" __global__ void __launch_bounds__(32) coal_rj3_4x4(const __grid_constant__ StkArgs a) { rj3_body<4, 4>(a); }
extern "C" __global__ void __launch_bounds__(32) coal_rj3_1x4(const __grid_constant__ StkArgs a) { rj3_body<1, 4>(a); }
extern "C" __global__ void __launch_bounds__(32) coal_rj3_8x2(const __grid_constant__ StkArgs a) { rj3_body<8, 2>(a); }

// G = 8 lanes per chain is the default (`rj`): at 4096 chains per GPU it is the best trade between lanes that idle
// when n+1 < G and warps per scheduler (B200: G=32 874 M, G=16 1168 M, G=8 1285 M, G=4 1000 M transitions/s;
// profiles/r01_coal_rj.md).
extern "C" __global__ void __launch_bounds__(BLOCK) coal_rj8(const __grid_constant__ StkArgs a) { rj_body<8>(a); }
extern "C" __global__ void __launch_bounds__(BLOCK) coal_rj32(const __grid_constant__ StkArgs a) { rj_body<32>(a); }
extern "C" __global__ void __launch_bounds__(BLOCK) coal_rj16(const __grid_constant__ StkArgs a) { rj_body<16>(a); }
extern "C" __global__ void __launch_bounds__(BLOCK) coal_rj4(const __grid_constant__ StkArgs a) { rj_body<4>(a); }

// joint log-density of every chain's current state (stk_logp; `lpdf`, Program.hs:787-812)
extern "C" __global__ void __launch_bounds__(BLOCK) coal_logp(const __grid_constant__ StkArgs a) {
  extern __shared__ __align__(16) double smem[];
  Chain<32> ch; ch.setup(a, smem);
  __syncthreads();
  const int n = ch.load(a);
  View v; v.set(ch.s, ch.g, 0, 0, 0.0, 0.0);
  const double lp = eval_f<32>(v, n, ch.ysm, ch.T, ch.k, ch.l, ch.warp_rounds(n));
  if (ch.active && ch.l == 0) a.out_lp[ch.chain] = lp;
}

// forward simulation of coalPrior (app/rj/Main.hs:115-122): n ~ truncated 1 nMax (poisson lam) by
// rejection (CC.hs:282-286), s = every 2nd of 2n+1 sorted uniform 0 t draws (orderedSample,
// Program.hs:1160-1165 sorts), g_j ~ gamma a b.  One sequential Philox stream per chain (transition
// counter 2^64-1), drawn by the group's lane 0.
extern "C" __global__ void __launch_bounds__(BLOCK) coal_simulate(const __grid_constant__ StkArgs a) {
  extern __shared__ __align__(16) double smem[];
  Chain<32> ch; ch.setup(a, smem);
  __syncthreads();
  int n = 1;
  if (ch.l == 0) {
    Rng rng; rng.init(a.seed, (u64)(a.chain0 + ch.cidx), ~0ULL);
    for (int it = 0; it < 1000; ++it) { n = draw_poisson(rng, (double)ch.lam); if (n >= 1 && n <= ch.nmax) break; }
    n = max(1, min(n, ch.nmax));
    double* buf = ch.s;                         // 64 contiguous doubles: s row then g row
    const int m = 2 * n + 1;
    for (int i = 0; i < m; ++i) {               // insertion sort as the draws arrive
      const double u = 0.0 + (ch.k.t - 0.0) * rng.u01();
      int p = i;
      while (p > 0 && buf[p - 1] > u) { buf[p] = buf[p - 1]; --p; }
      buf[p] = u;
    }
    for (int i = 1; i <= n; ++i) buf[i] = buf[2 * i - 1];   // s'!(2i), 1-based
    ch.s[0] = 0.0; ch.s[n + 1] = ch.k.t;
    for (int i = n + 2; i < ROW; ++i) ch.s[i] = 0.0;
    ch.g[0] = 0.0;
    for (int i = 1; i < ROW; ++i) ch.g[i] = i <= n + 1 ? draw_gamma(rng, ch.k.a, ch.k.b) : 0.0;
  }
  __syncwarp();
  n = __shfl_sync(0xffffffffu, n, 0);
  View v; v.set(ch.s, ch.g, 0, 0, 0.0, 0.0);
  const double lp = eval_f<32>(v, n, ch.ysm, ch.T, ch.k, ch.l, ch.warp_rounds(n));
  ch.store(a, n, lp);
}

// descriptor ------------------------------------------------------------------
// smem_fixed = (TAB_DOUBLES + chains_per_block * 64) * 8 + 16; + 8 bytes per element of y
static_assert(TAB_DOUBLES == 322, "TAB_DOUBLES");
extern "C" __device__ const char stk_desc[] =
    "stk-module 1\n"
    "name coal\n"
    "nparam 0\n"
    "field n state i32 1\n"
    "field s state f64 30\n"
    "field g state f64 31\n"
    "field lp state f64 1\n"
    "field y data f64 0 sorted\n"
    "param t 40907\n"
    "param lam 3\n"
    "param nmax 30\n"
    "param a 1\n"
    "param b 200\n"
    "record n s g\n"
    "kernel rj coal_rj2_4 32 4 17276 - 0 pre coal_draws 256 8\n"
    "kernel rj8 coal_rj8 128 16 10784 y 8\n"
    "kernel rj1 coal_rj1 32 32 53044 - 0\n"
    "kernel rj2 coal_rj2_8 32 8 23740 - 0 pre coal_draws 256 8\n"
    "kernel rj2x2 coal_rj2_2 32 2 14044 - 0 pre coal_draws 256 8\n"
    "kernel rj2x1 coal_rj2_1 32 1 12428 - 0 pre coal_draws 256 8\n"
    "kernel rj2x16 coal_rj2_16 32 16 36668 - 0 pre coal_draws 256 8\n"
    "kernel rj2x32 coal_rj2_32 32 32 62524 - 0 pre coal_draws 256 8\n"
    "kernel rj3_4x2 coal_rj3_4x2 32 8 23740 - 0 pre coal_draws 256 8\n"
    "kernel rj3_2x2 coal_rj3_2x2 32 4 17276 - 0 pre coal_draws 256 8\n"
    "kernel rj3_2x4 coal_rj3_2x4 32 8 23740 - 0 pre coal_draws 256 8\n"
    "kernel rj3_4x4 coal_rj3_4x4 32 16 36668 - 0 pre coal_draws 256 8\n"
    "kernel rj3_1x4 coal_rj3_1x4 32 4 17276 - 0 pre coal_draws 256 8\n"
    "kernel rj3_8x2 coal_rj3_8x2 32 16 36668 - 0 pre coal_draws 256 8\n"
    "kernel rj32 coal_rj32 128 4 4640 y 8\n"
    "kernel rj16 coal_rj16 128 8 6688 y 8\n"
    "kernel rj4 coal_rj4 128 32 18976 y 8\n"
    "entry logp coal_logp 128 4 4640 y 8\n"
    "entry simulate coal_simulate 128 4 4640 y 8\n";
