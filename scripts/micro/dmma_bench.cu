// Microbenchmark (B200): fp64 throughput of mma.sync.m8n8k4.f64 (DMMA) against plain DFMA, register operands only.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_bench dmma_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) k_dmma(double* out, int iters) {
  double c[16][2];
  for (int i = 0; i < 16; ++i) { c[i][0] = 0; c[i][1] = 0; }
  double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0; for (int i = 0; i < 16; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters) {
  double c[32];
  for (int i = 0; i < 32; ++i) c[i] = i;
  double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9 * threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 32; ++i) c[i] = fma(a, c[i], b);
  }
  double s = 0; for (int i = 0; i < 32; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  double* out; cudaMalloc(&out, 148 * 4 * 256 * 8);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int blocks_per_sm = 1; blocks_per_sm <= 4; blocks_per_sm *= 2) {
    const int grid = 148 * blocks_per_sm, iters = 20000;
    float ms;
    k_dmma<<<grid, 256>>>(out, 100); cudaDeviceSynchronize();
    cudaEventRecord(e0); k_dmma<<<grid, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    double fl = (double)grid * 8 /*warps*/ * iters * 16 * (2.0 * 8 * 8 * 4);
    printf("DMMA m8n8k4: %d CTA/SM x 256 thr: %.2f TFLOP/s (%.3f ms)\n", blocks_per_sm, fl / ms / 1e9, ms);
    k_dfma<<<grid, 256>>>(out, 100); cudaDeviceSynchronize();
    cudaEventRecord(e0); k_dfma<<<grid, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    fl = (double)grid * 256 * iters * 32 * 2.0;
    printf("DFMA       : %d CTA/SM x 256 thr: %.2f TFLOP/s (%.3f ms)\n", blocks_per_sm, fl / ms / 1e9, ms);
  }
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
