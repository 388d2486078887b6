// dp.cu — hand-tuned module for config C5: Dirichlet-process mixture  dpmm n  (dp.ipynb cells 1-3,
// app/dp/Main.hs:8-31):  sticks ~ beta 1 alpha, probs by stick breaking, atoms ~ uniform 0 100,
// values_j ~ normal atoms[stick_j] 1,  alpha = 5.
//   * `simulate`: the forward model truncated at K sticks — what the reference does with this model (Church backend).
//   * `sweep`: blocked Gibbs for the truncated stick-breaking representation (Ishwaran & James 2001); the reference
//     has no inference for this model (parity unpinned; the oracle is oracle/stk_oracle_dp.cpp).
//
// Pattern ("per-observation latent indices, two kernels per transition"):
//   * phase kernel `dp_assign` on a (32-chain groups) x (tiles of 2048 observations) grid: a LANE IS A CHAIN, so the
//     state arrays stored [element][chain] are read and the assignments z[i][chain] written as full 128-byte lines
//     (the one stream that must reach HBM: 4 N bytes per chain and sweep); y_i is a broadcast load that stays in L2
//     (shared by all chains); the chain's 32 log-weights and atoms sit in shared memory [k][lane], conflict-free;
//   * sufficient statistics are exact integers (n_k, sum_k round(y_i 2^36)) added with shared-memory then global
//     integer atomics: order-independent, hence bitwise reproducible and identical to the oracle's loop;
//   * main kernel `dp_params`: one warp per chain, lane k = stick k: suffix counts by shuffle scan, the two gamma
//     draws of the stick's beta posterior, the atom's truncated-normal posterior; resets the accumulators.
// Philox: counter (chain, sweep, slot): slot i/2 component i%2 -> u_i;  2^30 + 64k (+32) -> stick k;  2^30 + 2^16 + 64k -> atom k.
#include "../stk_device.cuh"
#include "../stk_fastmath.cuh"

using namespace stk;

namespace {

constexpr int KMAX = 32;
constexpr int TP = 2048;              // observations per tile (8 warps x 128 pairs)
constexpr u32 SB = 1u << 30, AB = (1u << 30) + (1u << 16);
constexpr double FIX = 68719476736.0; // 2^36

enum { F_V = 0, F_MU = 1, F_Z = 2, F_ACCN = 3, F_ACCS = 4, F_YSIM = 5, F_Y = 6 };
enum { PAR_ALPHA = 0, PAR_K = 1, PAR_LO = 2, PAR_HI = 3 };

// categorical draw over K <= 32 clusters with log-weights lpi_k - (y - mu_k)^2 / 2; tables are [k][32 lanes].
// Weights more than e^-50 below the largest are exactly zero (the oracle's rule), so only a few clusters per
// observation need an exp — but WHICH ones differs from lane to lane (a lane is a chain).  Looping over k with a
// branch around the exp makes the warp run all 32 exps (ncu r01dp: 2,120 instructions per warp-observation);
// instead each lane walks ITS OWN candidate bitmask, so the warp runs max-over-lanes(#candidates) ~ 3 iterations of
// one exp.  Sums run over the candidates in increasing k: identical to the full loop, which only adds zeros.
__device__ __forceinline__ int assign(double y, const double* __restrict__ lpi, const double* __restrict__ mu,
                                      const double* __restrict__ fm, double* __restrict__ wbuf, int lane, int K, double u) {
  double lw[KMAX];
  double m = neg_inf();
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    const double d = y - mu[k * 32 + lane];
    lw[k] = fma(-0.5 * d, d, lpi[k * 32 + lane]);      // clusters >= K carry lpi = -inf
    m = lw[k] > m ? lw[k] : m;
  }
  u32 mask = 0u;
#pragma unroll
  for (int k = 0; k < KMAX; ++k) mask |= (lw[k] - m < -50.0) ? 0u : (1u << k);
  // candidate weights go through a per-thread column of shared memory ([slot][thread]): a dynamic index costs one
  // STS/LDS there, where a register array would cost a select chain per slot
  double S = 0.0;
  int nc = 0;
  for (u32 mm = mask; mm; mm &= mm - 1u) {
    const int k = __ffs((int)mm) - 1;
    const double d = y - mu[k * 32 + lane];
    const double w = fast_exp(fma(-0.5 * d, d, lpi[k * 32 + lane]) - m, fm);   // table exp: 15 instructions, <= 1 ulp
    S += w;
    wbuf[nc * 256] = w;
    ++nc;
  }
  const double target = u * S;
  double cum = 0.0; int z = K, j = 0;
  for (u32 mm = mask; mm; mm &= mm - 1u, ++j) {
    cum += wbuf[j * 256];
    if (target < cum) { z = __ffs((int)mm); break; }
  }
  return min(z, K);
}

// log probs of the sticks of this lane's chain: lpi[k][lane] = log(v_k prod_{l<k} (1 - v_l)), v_K = 1
__device__ __forceinline__ void stage_tables(const StkArgs& a, long long cidx, int lane, int K, double* lpi, double* mus) {
  const double* V = (const double*)a.field[F_V];
  const double* M = (const double*)a.field[F_MU];
  double rem = 1.0;
  for (int k = 0; k < KMAX; ++k) {
    if (k < K) {
      const double vk = (k == K - 1) ? 1.0 : V[(long long)k * a.n_chains + cidx];
      lpi[k * 32 + lane] = log(vk * rem);
      rem *= (1.0 - vk);
      mus[k * 32 + lane] = M[(long long)k * a.n_chains + cidx];
    } else { lpi[k * 32 + lane] = neg_inf(); mus[k * 32 + lane] = 0.0; }
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
// assignment step: grid.x = 32-chain groups, grid.y = tiles of TP observations; block 256
// ---------------------------------------------------------------------------------------------------------
extern "C" __global__ void __launch_bounds__(256) dp_assign(const __grid_constant__ StkArgs a) {
  extern __shared__ __align__(16) double wcols[];      // [KMAX][256]: candidate weights of every thread
  __shared__ double lpi[KMAX * 32], mus[KMAX * 32];
  __shared__ unsigned cnt[KMAX * 32];
  __shared__ unsigned long long sum[KMAX * 32];
  __shared__ __align__(16) double fm[kFmTableDoubles];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int K = min((int)a.par[PAR_K], KMAX);
  const long long chain = (long long)blockIdx.x * 32 + lane;
  const bool active = chain < a.n_chains;
  const long long cidx = active ? chain : a.n_chains - 1;
  const long long N = a.flen[F_Y];
  for (int i = threadIdx.x; i < KMAX * 32; i += blockDim.x) { cnt[i] = 0u; sum[i] = 0ull; }
  fm_build_tables(fm, threadIdx.x, blockDim.x);
  if (w == 0) stage_tables(a, cidx, lane, K, lpi, mus);
  __syncthreads();
  const double* Y = (const double*)a.field[F_Y];
  int* Z = (int*)a.field[F_Z];
  const long long i0 = (long long)blockIdx.y * TP, i1 = min(i0 + TP, N);
  Rng rng; rng.init(a.seed, (u64)(a.chain0 + cidx), a.step0);
  for (long long i = i0 + 2 * w; i < i1; i += 16) {
    Rng r = rng; r.slot = (u32)(i >> 1);
    double u0, u1; r.u01x2(u0, u1);
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      if (i + e < i1) {
        const double y = __ldg(Y + i + e);
        const int z = assign(y, lpi, mus, fm, wcols + threadIdx.x, lane, K, e ? u1 : u0);
        if (active) {
          Z[(i + e) * a.n_chains + chain] = z;
          atomicAdd(&cnt[(z - 1) * 32 + lane], 1u);
          atomicAdd(&sum[(z - 1) * 32 + lane], (unsigned long long)__double2ll_rn(y * FIX));
        }
      }
    }
  }
  __syncthreads();
  int* AN = (int*)a.field[F_ACCN];
  unsigned long long* AS = (unsigned long long*)a.field[F_ACCS];
  for (int idx = threadIdx.x; idx < KMAX * 32; idx += blockDim.x) {
    const int k = idx >> 5, l = idx & 31;
    const long long c = (long long)blockIdx.x * 32 + l;
    if (c < a.n_chains && cnt[idx]) {
      atomicAdd(&AN[(long long)k * a.n_chains + c], (int)cnt[idx]);
      atomicAdd(&AS[(long long)k * a.n_chains + c], sum[idx]);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// sticks and atoms given the assignments: one warp per chain, lane k = stick k
// ---------------------------------------------------------------------------------------------------------
namespace {

__device__ __forceinline__ double draw_beta_stick(const Rng& rng, int k, double a, double b) {
  Rng g = rng; g.slot = SB + 64u * (u32)k;
  const double X = draw_gamma(g, a, 1.0);
  g.slot = SB + 64u * (u32)k + 32u;
  const double Y = draw_gamma(g, b, 1.0);
  return X / (X + Y);
}

}  // namespace

extern "C" __global__ void __launch_bounds__(128) dp_params(const __grid_constant__ StkArgs a) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int K = min((int)a.par[PAR_K], KMAX);
  const double alpha = a.par[PAR_ALPHA], lo = a.par[PAR_LO], hi = a.par[PAR_HI];
  const long long chain = (long long)blockIdx.x * 4 + w;
  if (chain >= a.n_chains) return;
  int* AN = (int*)a.field[F_ACCN];
  long long* AS = (long long*)a.field[F_ACCS];
  double* V = (double*)a.field[F_V];
  double* M = (double*)a.field[F_MU];
  const long long off = (long long)lane * a.n_chains + chain;
  const int n = lane < K ? AN[off] : 0;
  const long long s = lane < K ? AS[off] : 0;
  AN[off] = 0; AS[off] = 0;                             // ready for the next sweep's assignment step
  int incl = n;                                         // inclusive prefix sum over sticks
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  const int after = total - incl;                       // sum_{l > k} n_l
  Rng rng; rng.init(a.seed, (u64)(a.chain0 + chain), a.step0);
  double v = 1.0, mu = 0.0;
  if (lane < K) {
    if (lane < K - 1) v = draw_beta_stick(rng, lane, 1.0 + (double)n, alpha + (double)after);
    Rng g = rng; g.slot = AB + 64u * (u32)lane;
    if (n > 0) {
      const double mean = ((double)s / FIX) / (double)n, sd = 1.0 / sqrt((double)n);
      double m = mean;
      for (int it = 0; it < 32; ++it) { m = mean + sd * g.normal(); if (m > lo && m < hi) break; }
      mu = fmin(fmax(m, lo), hi);
    } else mu = lo + (hi - lo) * g.u01();
    V[off] = v; M[off] = mu;
  }
  long long slot;
  if (record_slot(a, 0, &slot)) {
    record_value(a, slot, lane, chain, lane < K ? v : 0.0);
    record_value(a, slot, KMAX + lane, chain, lane < K ? mu : 0.0);
  }
}

// ---------------------------------------------------------------------------------------------------------
// forward simulation of dpmm n truncated at K sticks (sweep counter 2^64-1): dp_sim_points on the 2-D grid draws
// stick_j and values_j for every observation; dp_sim_params stores the sticks and atoms of the same Philox draws
// ---------------------------------------------------------------------------------------------------------
namespace {
__device__ __forceinline__ void prior_params(const StkArgs& a, long long gchain, int k, int K, double* v, double* mu) {
  Rng rng; rng.init(a.seed, (u64)gchain, ~0ULL);
  const double vv = draw_beta_stick(rng, k, 1.0, a.par[PAR_ALPHA]);
  *v = (k == K - 1) ? 1.0 : vv;
  Rng g = rng; g.slot = AB + 64u * (u32)k;
  *mu = a.par[PAR_LO] + (a.par[PAR_HI] - a.par[PAR_LO]) * g.u01();
}
}  // namespace

extern "C" __global__ void __launch_bounds__(256) dp_sim_points(const __grid_constant__ StkArgs a) {
  __shared__ double pr[KMAX * 32], mus[KMAX * 32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int K = min((int)a.par[PAR_K], KMAX);
  const long long chain = (long long)blockIdx.x * 32 + lane;
  const bool active = chain < a.n_chains;
  const long long cidx = active ? chain : a.n_chains - 1;
  const long long N = a.flen[F_Y];
  for (int k = w; k < KMAX; k += 8) {
    double v = 0.0, mu = 0.0;
    if (k < K) prior_params(a, a.chain0 + cidx, k, K, &v, &mu);
    pr[k * 32 + lane] = v; mus[k * 32 + lane] = mu;
  }
  __syncthreads();
  if (w == 0) {                                          // sticks -> probs (stick breaking, dp.ipynb cell 1)
    double rem = 1.0;
    for (int k = 0; k < KMAX; ++k) { const double v = pr[k * 32 + lane]; pr[k * 32 + lane] = v * rem; rem *= (1.0 - v); }
  }
  __syncthreads();
  int* Z = (int*)a.field[F_Z];
  double* YS = (double*)a.field[F_YSIM];
  const long long i0 = (long long)blockIdx.y * TP, i1 = min(i0 + TP, N);
  Rng rng; rng.init(a.seed, (u64)(a.chain0 + cidx), ~0ULL);
  for (long long i = i0 + w; i < i1; i += 8) {
    Rng r = rng; r.slot = (u32)(i >> 1);
    double u0, u1; r.u01x2(u0, u1);
    const double u = (i & 1) ? u1 : u0;
    double cum = 0.0; int z = K; bool found = false;
    for (int k = 0; k < K; ++k) { cum += pr[k * 32 + lane]; if (!found && u < cum) { z = k + 1; found = true; } }
    Rng rn = rng; rn.slot = (1u << 29) + (u32)i;
    const double val = mus[(z - 1) * 32 + lane] + rn.normal();
    if (active) { Z[i * a.n_chains + chain] = z; YS[i * a.n_chains + chain] = val; }
  }
}

extern "C" __global__ void __launch_bounds__(128) dp_sim_params(const __grid_constant__ StkArgs a) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int K = min((int)a.par[PAR_K], KMAX);
  const long long chain = (long long)blockIdx.x * 4 + w;
  if (chain >= a.n_chains || lane >= K) return;
  double v, mu;
  prior_params(a, a.chain0 + chain, lane, K, &v, &mu);
  ((double*)a.field[F_V])[(long long)lane * a.n_chains + chain] = v;
  ((double*)a.field[F_MU])[(long long)lane * a.n_chains + chain] = mu;
}

extern "C" __device__ const char stk_desc[] =
    "stk-module 1\n"
    "name dp\n"
    "nparam 0\n"
    "field v state f64 32\n"
    "field mu state f64 32\n"
    "field z state i32 @y\n"
    "field acc_n state i32 32\n"
    "field acc_s state f64 32\n"
    "field ysim state f64 @y\n"
    "field y data f64 0\n"
    "param alpha 5\n"
    "param K 32\n"
    "param lo 0\n"
    "param hi 100\n"
    "record v mu\n"
    "kernel sweep dp_params 128 4 0 - 0 phase dp_assign 256 32 y 2048 65536\n"
    "entry simulate dp_sim_params 128 4 0 - 0 phase dp_sim_points 256 32 y 2048 0\n";
