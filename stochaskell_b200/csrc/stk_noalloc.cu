// stk_noalloc.cu — the one entry point of libstk_cuda.so that allocates nothing: stk_diag_partial_noalloc.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
//
// It is the variant of stk_diag_partial's two kernels (models/diag.cu: diag_moments, diag_acov) whose per-chain scratch — the
// [P][C] table of chain means the first kernel writes and the second reads — lives in a file-scope __device__ array of
// STK_NOALLOC_MAX_MEANS doubles (4 MiB, held by every GPU that loads this library until the process exits) instead of a
// cudaMalloc'ed buffer.  The caller passes the draws and the output buffer as DEVICE pointers and the stream; the call
// enqueues a memset of the output and two kernels on that stream and returns without synchronising.
//
// Sharing: every call on a device uses the same array, and it keeps the last call's contents.  Nothing in it has to start at
// zero (the first kernel overwrites every entry the second one reads); the OUTPUT buffer does, and is cleared here.  Two calls
// that overlap on the GPU (different streams, no dependency between them) race on the array; serialise them.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <mutex>

#include "../../include/stk_cuda.h"
#include "stk_device.cuh"

extern "C" int stk_fail_(int code, const char* msg);      // stk_runtime.cpp: sets stk_last_error()
extern "C" void stk_count_launches_(int n);               // stk_runtime.cpp: stk_launch_count()

#define STK_NOALLOC_MAX_MEANS (512 * 1024)                // P * C <= 524288 doubles = 4 MiB per GPU

static __device__ __align__(256) double g_noalloc_means[STK_NOALLOC_MAX_MEANS];

namespace {

// per (value r, chain c): mean and unbiased variance over the N recorded draws; the block's sums of
// (1, mean, mean^2, var) are added to buf[0..3][r] — the same arithmetic as diag_moments, per chain, so that the two
// entry points differ only in the order of the atomic additions
__global__ void __launch_bounds__(256)
noalloc_moments(const double* __restrict__ draws, long long N, int P, long long C, double* __restrict__ means,
                double* __restrict__ buf) {
  const int r = blockIdx.y;
  const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  double q[4] = {0.0, 0.0, 0.0, 0.0};
  if (c < C) {
    double s = 0.0;
    for (long long i = 0; i < N; ++i) s += draws[(i * P + r) * C + c];
    const double m = s / (double)N;
    double s2 = 0.0;
    for (long long i = 0; i < N; ++i) { const double d = draws[(i * P + r) * C + c] - m; s2 = fma(d, d, s2); }
    means[(long long)r * C + c] = m;
    q[0] = 1.0; q[1] = m; q[2] = m * m; q[3] = N > 1 ? s2 / (double)(N - 1) : 0.0;
  }
  __shared__ double part[4][8];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const double x = stk::group_sum<32>(q[k]);
    if ((threadIdx.x & 31) == 0) part[k][threadIdx.x >> 5] = x;
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double x = 0.0;
    for (int w = 0; w < 8; ++w) x += part[threadIdx.x][w];
    atomicAdd(&buf[(long long)threadIdx.x * P + r], x);
  }
}

// autocovariance sums of one (group of CB chains, value r): the centred series staged in shared memory [N][CB], thread
// (lag slice, chain) accumulates acov_t = sum_i x_i x_{i+t} / (N - t), summed over the group's chains by shuffles and added
// to acov[r][t]
__global__ void __launch_bounds__(256)
noalloc_acov(const double* __restrict__ draws, const double* __restrict__ means, long long N, int P, long long C, int CB,
             long long max_lag, double* __restrict__ acov) {
  extern __shared__ double series[];
  const int r = blockIdx.y;
  const long long c0 = (long long)blockIdx.x * CB;
  for (long long idx = threadIdx.x; idx < N * CB; idx += blockDim.x) {
    const long long i = idx / CB; const long long c = c0 + (idx - i * CB);
    series[idx] = c < C ? draws[(i * P + r) * C + c] - means[(long long)r * C + c] : 0.0;
  }
  __syncthreads();
  const int k = threadIdx.x % CB, slice = threadIdx.x / CB, S = blockDim.x / CB;
  for (long long t0 = 0; t0 <= max_lag; t0 += S) {          // uniform trip count: the whole warp reaches the shuffles
    const long long t = t0 + slice;
    const bool valid = t <= max_lag && t < N;
    double s = 0.0;
    if (valid) {
      for (long long i = 0; i + t < N; ++i) s = fma(series[i * CB + k], series[(i + t) * CB + k], s);
      s /= (double)(N - t);
    }
    for (int o = CB / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (k == 0 && valid) atomicAdd(&acov[(long long)r * (max_lag + 1) + t], s);
  }
}

constexpr int kMaxDevices = 64;
constexpr size_t kMaxSeriesBytes = 200 * 1024;            // the shared-memory series of noalloc_acov, as stk_diag_partial sizes it
std::mutex g_setup_mutex;
double* g_means_ptr[kMaxDevices] = {};                    // DEVICE address of g_noalloc_means on each device (the host-side
                                                          // name `g_noalloc_means` is a stand-in, never passed to a kernel)

int failf(int code, const char* fmt, long long a = 0, long long b = 0, long long c = 0) {
  char msg[512];
  snprintf(msg, sizeof msg, fmt, a, b, c);
  return stk_fail_(code, msg);
}
int cuda_fail(const char* what, cudaError_t e) {
  char msg[512];
  snprintf(msg, sizeof msg, "stk_diag_partial_noalloc: %s: %s", what, cudaGetErrorString(e));
  return stk_fail_(STK_ERR_CUDA, msg);
}

// once per device, with that device current: the array's device address and the kernel's shared-memory ceiling
int setup_device(int dev, double** means) {
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (!g_means_ptr[dev]) {
    double* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_means);
    if (e != cudaSuccess) return cuda_fail("cudaGetSymbolAddress", e);
    e = cudaFuncSetAttribute((const void*)noalloc_acov, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSeriesBytes);
    if (e != cudaSuccess) return cuda_fail("cudaFuncSetAttribute", e);
    g_means_ptr[dev] = p;
  }
  *means = g_means_ptr[dev];
  return STK_OK;
}

}  // namespace

extern "C" int stk_diag_partial_noalloc(const double* d_draws, int64_t n_draws, int n_values, int64_t n_chains, int64_t max_lag,
                                        double* d_buf, size_t bytes, stk_stream stream) {
  if (!d_draws || !d_buf) return stk_fail_(STK_ERR_ARG, "stk_diag_partial_noalloc: null device pointer");
  if (n_draws < 2 || n_values < 1 || n_chains < 1) return stk_fail_(STK_ERR_ARG, "stk_diag_partial_noalloc: need n_draws >= 2, n_values >= 1, n_chains >= 1");
  const long long N = n_draws, C = n_chains; const int P = n_values;
  if ((long long)P > STK_NOALLOC_MAX_MEANS / C)
    return failf(STK_ERR_UNSUPPORTED, "stk_diag_partial_noalloc: n_values * n_chains = %lld exceeds the static scratch of %lld doubles (4 MiB); use stk_diag_partial",
                 (long long)P * C, (long long)STK_NOALLOC_MAX_MEANS);
  if (P > 65535) return stk_fail_(STK_ERR_UNSUPPORTED, "stk_diag_partial_noalloc: more than 65535 values per draw");
  long long ml = (max_lag < 0 || max_lag > N - 1) ? N - 1 : max_lag;
  const size_t want = sizeof(double) * (size_t)P * (size_t)(4 + ml + 1);
  if (bytes != want) return failf(STK_ERR_ARG, "stk_diag_partial_noalloc: output buffer must be %lld bytes, got %lld", (long long)want, (long long)bytes);
  int CB = 32;
  while (CB > 1 && (size_t)N * CB * 8 > kMaxSeriesBytes) CB >>= 1;
  const size_t smem = (size_t)N * CB * 8;
  if (smem > kMaxSeriesBytes) return failf(STK_ERR_UNSUPPORTED, "stk_diag_partial_noalloc: too many draws (%lld, at most %lld); thin the record", N, (long long)(kMaxSeriesBytes / 8));

  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail("cudaGetDevice", e);
  if (dev < 0 || dev >= kMaxDevices) return stk_fail_(STK_ERR_UNSUPPORTED, "stk_diag_partial_noalloc: device ordinal out of range");
  double* means = nullptr;
  int rc = setup_device(dev, &means);
  if (rc) return rc;

  cudaStream_t st = (cudaStream_t)stream;
  e = cudaMemsetAsync(d_buf, 0, want, st);                  // the additive output starts at zero; the scratch need not
  if (e != cudaSuccess) return cuda_fail("cudaMemsetAsync", e);
  noalloc_moments<<<dim3((unsigned)((C + 255) / 256), (unsigned)P), 256, 0, st>>>(d_draws, N, P, C, means, d_buf);
  noalloc_acov<<<dim3((unsigned)((C + CB - 1) / CB), (unsigned)P), 256, smem, st>>>(d_draws, means, N, P, C, CB, ml, d_buf + 4 * (size_t)P);
  stk_count_launches_(2);
  e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail("kernel launch", e);
  return STK_OK;
}
