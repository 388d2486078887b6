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
    "kernel rj coal_rj8 128 16 10784 y 8\n"
    "kernel rj32 coal_rj32 128 4 4640 y 8\n"
    "kernel rj16 coal_rj16 128 8 6688 y 8\n"
    "kernel rj4 coal_rj4 128 32 18976 y 8\n"
    "entry logp coal_logp 128 4 4640 y 8\n"
    "entry simulate coal_simulate 128 4 4640 y 8\n";
