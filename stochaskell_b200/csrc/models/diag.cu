// diag.cu — cross-chain diagnostics over the recorded draws ring
//   draws[slot][r][chain]  (slot-major, chain-minor; see stk_abi.h)
// The reference computes no diagnostics itself (CmdStan's stansummary
// exists but is never called); the formulas follow the vendored Stan so that a
// user of hmcStan gets the numbers stansummary would print:
//   cmdstan/stan/src/stan/analyze/mcmc/compute_effective_sample_size.hpp:30-118
//   cmdstan/stan/src/stan/analyze/mcmc/compute_potential_scale_reduction.hpp
//   stan_math .../prim/mat/fun/autocovariance.hpp:64-68 (acov_t = sum c_i c_{i+t} / (N-t))
// Every output is ADDITIVE over chains, so ranks all-reduce the buffer (NCCL,
// SURVEY 8e) and finish on the host with stk_diag_finish.
#include "../stk_device.cuh"

// per (r, chain): mean and unbiased variance over the N recorded draws, then
// block-reduce over chains into buf[0..3][r] = (count, sum mean, sum mean^2, sum var)
extern "C" __global__ void __launch_bounds__(256)
diag_moments(const double* __restrict__ draws, long long N, int nrec, long long C,
             double* __restrict__ chain_mean /*[nrec][C]*/, double* __restrict__ buf /*[4][nrec]*/) {
  const int r = blockIdx.y;
  const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  double m = 0.0, v = 0.0, cnt = 0.0;
  if (c < C) {
    double s = 0.0;
    for (long long i = 0; i < N; ++i) s += draws[(i * nrec + r) * C + c];
    m = s / (double)N;
    double s2 = 0.0;
    for (long long i = 0; i < N; ++i) { double d = draws[(i * nrec + r) * C + c] - m; s2 = fma(d, d, s2); }
    v = N > 1 ? s2 / (double)(N - 1) : 0.0;
    chain_mean[(long long)r * C + c] = m;
    cnt = 1.0;
  }
  __shared__ double sh[4][8];
  double q[4] = {cnt, m, m * m, v};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    double x = stk::group_sum<32>(q[k]);
    if ((threadIdx.x & 31) == 0) sh[k][threadIdx.x >> 5] = x;
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double x = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) x += sh[threadIdx.x][w];
    atomicAdd(&buf[(long long)threadIdx.x * nrec + r], x);
  }
}

// autocovariance sums: block = (chain group of CB chains, r); the group's
// centred series is staged in shared memory [N][CB]; thread (lag slice, chain)
// accumulates acov_t for t = slice, slice+S, ...; the per-chain values are
// summed over the group's chains and atomically added to acov[r][t].
extern "C" __global__ void __launch_bounds__(256)
diag_acov(const double* __restrict__ draws, const double* __restrict__ chain_mean,
          long long N, int nrec, long long C, int CB, long long max_lag,
          double* __restrict__ acov /*[nrec][max_lag+1]*/) {
  extern __shared__ double ser[];  // [N][CB]
  const int r = blockIdx.y;
  const long long c0 = (long long)blockIdx.x * CB;
  for (long long idx = threadIdx.x; idx < N * CB; idx += blockDim.x) {
    long long i = idx / CB; int k = (int)(idx - i * CB);
    long long c = c0 + k;
    ser[idx] = c < C ? draws[(i * nrec + r) * C + c] - chain_mean[(long long)r * C + c] : 0.0;
  }
  __syncthreads();
  const int k = threadIdx.x % CB;          // chain within group (fastest: conflict-free smem reads)
  const int slice = threadIdx.x / CB;
  const int S = blockDim.x / CB;
  // uniform trip count: every thread of a warp reaches the shuffles together
  for (long long t0 = 0; t0 <= max_lag; t0 += S) {
    const long long t = t0 + slice;
    const bool valid = t <= max_lag && t < N;
    double s = 0.0;
    if (valid) {
      for (long long i = 0; i + t < N; ++i) s = fma(ser[i * CB + k], ser[(i + t) * CB + k], s);
      s /= (double)(N - t);
    }
    // sum over the CB chains of the group (CB is a power of two <= 32, lanes contiguous)
    for (int o = CB / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (k == 0 && valid) atomicAdd(&acov[(long long)r * (max_lag + 1) + t], s);
  }
}

extern "C" __device__ const char stk_desc[] =
    "stk-module 1\n"
    "name diag\n"
    "nparam 0\n";
