#include <cstdio>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  unsigned long long ra, rb, rc, rd;
  ra = *reinterpret_cast<unsigned long long*>(&a);
  rb = *reinterpret_cast<unsigned long long*>(&b);
  rc = *reinterpret_cast<unsigned long long*>(&c);
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  return *reinterpret_cast<float2*>(&rd);
}
template <int MODE>
__global__ void k(float* out, int iters, float s) {
  float2 acc[16];
  for (int i = 0; i < 16; i++) acc[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
  float2 a = make_float2(s, s * 1.0001f), b = make_float2(0.9999f, 1.00001f);
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
      if (MODE == 0) {
        acc[i].x = fmaf(a.x, b.x, acc[i].x);
        acc[i].y = fmaf(a.y, b.y, acc[i].y);
      } else {
        acc[i] = ffma2(a, b, acc[i]);
      }
    }
  }
  float r = 0;
  for (int i = 0; i < 16; i++) r += acc[i].x + acc[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
int main() {
  float* d;
  cudaMalloc(&d, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int mode = 0; mode < 2; mode++) {
    for (int rep = 0; rep < 2; rep++) {
      cudaEventRecord(e0);
      if (mode == 0) k<0><<<148 * 4, 256>>>(d, iters, 1.0f); else k<1><<<148 * 4, 256>>>(d, iters, 1.0f);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double fma = 148.0 * 4 * 256 * (double)iters * 32;
      printf("mode %d: %.3f ms  %.1f TFLOP/s fp32\n", mode, ms, 2 * fma / ms / 1e9);
    }
  }
  return 0;
}
