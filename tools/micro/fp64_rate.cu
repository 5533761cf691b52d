// Microbenchmark: FP64 SIMT DFMA vs DMMA (mma.sync.m8n8k4.f64) throughput and dependent-issue latency on one B200.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void dfma_tp(double* out, int iters) {
  double a[8];
  for (int i = 0; i < 8; i++) a[i] = threadIdx.x * 1e-3 + i;
  const double b = 1.0000001, c = 1e-9;
  for (int it = 0; it < iters; it++)
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = fma(a[i], b, c);
  double s = 0; for (int i = 0; i < 8; i++) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void dfma_lat(double* out, int iters) {
  double a = threadIdx.x * 1e-3;
  const double b = 1.0000001, c = 1e-9;
  for (int it = 0; it < iters; it++) { a = fma(a, b, c); a = fma(a, b, c); a = fma(a, b, c); a = fma(a, b, c); }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a;
}
__global__ void dmma_tp(double* out, int iters) {
  double c0[2] = {0, 0}, c1[2] = {0, 0}, c2[2] = {0, 0}, c3[2] = {0, 0};
  double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6;
  for (int it = 0; it < iters; it++) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0[0]), "+d"(c0[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c1[0]), "+d"(c1[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c2[0]), "+d"(c2[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c3[0]), "+d"(c3[1]) : "d"(a), "d"(b));
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = c0[0] + c0[1] + c1[0] + c1[1] + c2[0] + c2[1] + c3[0] + c3[1];
}
__global__ void dmma_lat(double* out, int iters) {
  double c0[2] = {0, 0};
  double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6;
  for (int it = 0; it < iters; it++) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0[0]), "+d"(c0[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0[0]), "+d"(c0[1]) : "d"(a), "d"(b));
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = c0[0] + c0[1];
}
template <typename F> float timeit(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize(); cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}
int main() {
  double* out; cudaMalloc(&out, 148 * 8 * 1024 * 8);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const double ghz = clk * 1e-6;
  const int iters = 20000;
  for (int warps : {1, 4, 8, 16, 32}) {
    float ms = timeit([&] { dfma_tp<<<148, warps * 32>>>(out, iters); });
    double fma_per_clk_sm = (double)warps * 32 * 8 * iters / (ms * 1e-3 * ghz * 1e9);
    printf("DFMA  tp  warps/SM=%2d: %.3f ms  -> %.1f FMA/clk/SM (%.2f TFLOP/s chip)\n", warps, ms, fma_per_clk_sm, fma_per_clk_sm * 2 * 148 * ghz * 1e-3);
  }
  { float ms = timeit([&] { dfma_lat<<<1, 32>>>(out, iters); }); printf("DFMA  dependent latency: %.1f cycles\n", ms * 1e-3 * ghz * 1e9 / (4.0 * iters)); }
  for (int warps : {1, 4, 8, 16, 32}) {
    float ms = timeit([&] { dmma_tp<<<148, warps * 32>>>(out, iters); });
    double fma_per_clk_sm = (double)warps * 4 * 256 * iters / (ms * 1e-3 * ghz * 1e9);
    printf("DMMA  tp  warps/SM=%2d: %.3f ms  -> %.1f FMA/clk/SM (%.2f TFLOP/s chip)\n", warps, ms, fma_per_clk_sm, fma_per_clk_sm * 2 * 148 * ghz * 1e-3);
  }
  { float ms = timeit([&] { dmma_lat<<<1, 32>>>(out, iters); }); printf("DMMA  dependent latency: %.1f cycles\n", ms * 1e-3 * ghz * 1e9 / (2.0 * iters)); }
  printf("clock %.3f GHz\n", ghz);
  return 0;
}
