// Microbenchmark: latency of factoring one 32x32 fp64 SPD block by one warp / one CTA, several formulations.
// Timed with clock64 inside the kernel (single CTA), averaged over reps; checked against a host Cholesky.
#include <cstdio>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
constexpr int B = 32;
__device__ __forceinline__ double shfl_d(double v, int src) {
  int lo = __double2loint(v), hi = __double2hiint(v);
  lo = __shfl_sync(0xffffffffu, lo, src); hi = __shfl_sync(0xffffffffu, hi, src);
  return __hiloint2double(hi, lo);
}
__device__ __forceinline__ double fast_rsqrt(double x) {
  double y = (double)rsqrtf((float)x); const double hx = 0.5 * x;
  y = y * (1.5 - hx * y * y); y = y * (1.5 - hx * y * y); return y;
}
// V1: lane = row in registers, right-looking, shuffles broadcast the pivot column (fully unrolled)
__device__ void v1(double (*D)[B + 1], int lane) {
  double a[B];
#pragma unroll
  for (int c = 0; c < B; c++) a[c] = D[lane][c];
#pragma unroll
  for (int j = 0; j < B; j++) {
    const double djj = shfl_d(a[j], j);
    const double inv = fast_rsqrt(djj);
    const double l = (lane == j) ? djj * inv : a[j] * inv;
    a[j] = l;
#pragma unroll
    for (int c = j + 1; c < B; c++) { const double lc = shfl_d(l, c); if (lane >= c) a[c] -= l * lc; }
  }
#pragma unroll
  for (int c = 0; c < B; c++) D[lane][c] = (c <= lane) ? a[c] : 0.0;
}
// V2: rolled left-looking on shared memory
__device__ void v2(double (*D)[B + 1], int lane) {
  for (int j = 0; j < B; j++) {
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0; int k = 0;
    for (; k + 4 <= j; k += 4) { s0 = fma(D[lane][k], D[j][k], s0); s1 = fma(D[lane][k+1], D[j][k+1], s1); s2 = fma(D[lane][k+2], D[j][k+2], s2); s3 = fma(D[lane][k+3], D[j][k+3], s3); }
    for (; k < j; k++) s0 = fma(D[lane][k], D[j][k], s0);
    const double v = D[lane][j] - ((s0 + s1) + (s2 + s3));
    const double djj = shfl_d(v, j); const double inv = fast_rsqrt(djj);
    __syncwarp();
    if (lane >= j) D[lane][j] = (lane == j) ? djj * inv : v * inv;
    __syncwarp();
  }
}
// V3: lane = row in registers, right-looking, pivot column broadcast through shared memory (fully unrolled)
__device__ void v3(double (*D)[B + 1], double* col, int lane) {
  double a[B];
#pragma unroll
  for (int c = 0; c < B; c++) a[c] = D[lane][c];
#pragma unroll
  for (int j = 0; j < B; j++) {
    const double djj = shfl_d(a[j], j);
    const double inv = fast_rsqrt(djj);
    const double l = (lane == j) ? djj * inv : a[j] * inv;
    a[j] = l;
    col[lane] = l;
    __syncwarp();
#pragma unroll
    for (int c = j + 1; c < B; c++) a[c] = fma(-l, col[c], a[c]);  // rows < c compute garbage in their upper part: never read
    __syncwarp();
  }
#pragma unroll
  for (int c = 0; c < B; c++) D[lane][c] = (c <= lane) ? a[c] : 0.0;
}
// V4: as V3 but the pivot for step j+1 is updated first and its rsqrt overlaps the rest of the rank-1 update
__device__ void v4(double (*D)[B + 1], double* col, int lane) {
  double a[B];
#pragma unroll
  for (int c = 0; c < B; c++) a[c] = D[lane][c];
  double inv = fast_rsqrt(shfl_d(a[0], 0));
#pragma unroll
  for (int j = 0; j < B; j++) {
    const double l = a[j] * inv;  // lane j: a_jj * inv = sqrt(a_jj)
    a[j] = l;
    col[lane] = l;
    __syncwarp();
    double inv_next = 0.0;
    if (j + 1 < B) {
      a[j + 1] = fma(-l, col[j + 1], a[j + 1]);
      inv_next = fast_rsqrt(shfl_d(a[j + 1], j + 1));
    }
#pragma unroll
    for (int c = j + 2; c < B; c++) a[c] = fma(-l, col[c], a[c]);
    inv = inv_next;
    __syncwarp();
  }
#pragma unroll
  for (int c = 0; c < B; c++) D[lane][c] = (c <= lane) ? a[c] : 0.0;
}
// 1/x: fp32 reciprocal seed + two Newton steps in fp64 (two dependent FMAs each)
__device__ __forceinline__ double fast_rcp(double x) {
  double y = (double)__frcp_rn((float)x);
  double t = fma(-x, y, 1.0); y = fma(y, t, y);
  t = fma(-x, y, 1.0); y = fma(y, t, y);
  return y;
}
// V5: square-root-free pivot chain.  The rank-1 update uses the UNSCALED column (stored to shared memory before the
// reciprocal is known, so the broadcast overlaps the reciprocal) and m = a_j / d_jj; the rsqrt scaling of L is applied
// once at the end, off the chain.  Column buffer double-buffered: one __syncwarp per step.
__device__ void v5(double (*D)[B + 1], double (*col)[B], double* inv, int lane) {
  double a[B];
#pragma unroll
  for (int c = 0; c < B; c++) a[c] = D[lane][c];
#pragma unroll
  for (int j = 0; j < B; j++) {
    col[j & 1][lane] = a[j];
    const double djj = shfl_d(a[j], j);
    const double m = a[j] * fast_rcp(djj);
    __syncwarp();
#pragma unroll
    for (int c = j + 1; c < B; c++) a[c] = fma(-m, col[j & 1][c], a[c]);
  }
  double d = 0.0;
#pragma unroll
  for (int c = 0; c < B; c++) d = (c == lane) ? a[c] : d;
  inv[lane] = fast_rsqrt(d);
  __syncwarp();
#pragma unroll
  for (int c = 0; c < B; c++) D[lane][c] = (c <= lane) ? a[c] * inv[c] : 0.0;
}
// V6: as V5, the pivot broadcast through shared memory as well (no shuffles): every lane reads d_jj = col[j]
__device__ void v6(double (*D)[B + 1], double (*col)[B], double* inv, int lane) {
  double a[B];
#pragma unroll
  for (int c = 0; c < B; c++) a[c] = D[lane][c];
#pragma unroll
  for (int j = 0; j < B; j++) {
    col[j & 1][lane] = a[j];
    __syncwarp();
    const double m = a[j] * fast_rcp(col[j & 1][j]);
#pragma unroll
    for (int c = j + 1; c < B; c++) a[c] = fma(-m, col[j & 1][c], a[c]);
  }
  double d = 0.0;
#pragma unroll
  for (int c = 0; c < B; c++) d = (c == lane) ? a[c] : d;
  inv[lane] = fast_rsqrt(d);
  __syncwarp();
#pragma unroll
  for (int c = 0; c < B; c++) D[lane][c] = (c <= lane) ? a[c] * inv[c] : 0.0;
}
// V7: V3 folded: ONE unrolled 16-pivot body over 32 register columns, run twice (the second time on the columns 16..31
// moved down to registers 0..15, the upper 16 registers then hold don't-care values): 15 KB of code instead of 40 KB,
// 752 instead of 496 FMAs per lane.  col needs 48 entries (reads up to base + 31).
__device__ void v7(double (*D)[B + 1], double* col, int lane) {
  double a[B];
#pragma unroll
  for (int c = 0; c < B; c++) a[c] = D[lane][c];
#pragma unroll 1
  for (int base = 0; base < B; base += 16) {
#pragma unroll
    for (int j = 0; j < 16; j++) {
      const double djj = shfl_d(a[j], base + j);
      const double inv = fast_rsqrt(djj);
      const double l = (lane == base + j) ? djj * inv : a[j] * inv;
      a[j] = l;
      col[lane] = l;
      __syncwarp();
#pragma unroll
      for (int c = j + 1; c < B; c++) a[c] = fma(-l, col[base + c], a[c]);
      __syncwarp();
    }
#pragma unroll
    for (int c = 0; c < 16; c++) {
      D[lane][base + c] = (base + c <= lane) ? a[c] : 0.0;
      a[c] = a[c + 16];
    }
  }
}
template <int V>
__global__ void bench(const double* in, double* out, long long* cycles, int reps) {
  __shared__ double D[B][B + 1];
  __shared__ double col[B + 16];
  __shared__ double col2[2][B];
  __shared__ double Dx[B][B + 1];
  __shared__ double inv[B];
  const int lane = threadIdx.x;
  long long tot = 0;
  for (int r = 0; r < reps; r++) {
    for (int c = 0; c < B; c++) D[lane][c] = in[lane * B + c];
    if (V >= 10) {  // cold instruction cache: ~100 KB of other straight-line code between the timed calls
      for (int c = 0; c < B; c++) Dx[lane][c] = in[lane * B + c];
      __syncwarp();
      v1(Dx, lane);
      for (int c = 0; c < B; c++) Dx[lane][c] = in[lane * B + c];
      __syncwarp();
      v5(Dx, col2, inv, lane);
      for (int c = 0; c < B; c++) Dx[lane][c] = in[lane * B + c];
      __syncwarp();
      v4(Dx, col, lane);
      if (Dx[lane][0] == 123.456) D[lane][0] = 0.0;
    }
    __syncwarp();
    const long long t0 = clock64();
    if (V == 1) v1(D, lane);
    if (V == 2) v2(D, lane);
    if (V == 3) v3(D, col, lane);
    if (V == 4) v4(D, col, lane);
    if (V == 5) v5(D, col2, inv, lane);
    if (V == 6) v6(D, col2, inv, lane);
    if (V == 7) v7(D, col, lane);
    if (V == 13) v3(D, col, lane);
    if (V == 17) v7(D, col, lane);
    __syncwarp();
    tot += clock64() - t0;
  }
  for (int c = 0; c < B; c++) out[lane * B + c] = (c <= lane) ? D[lane][c] : 0.0;
  if (lane == 0) *cycles = tot / reps;
}
// V3 inside a 256-thread CTA with 200 KB of dynamic shared memory (the solver kernel's configuration): warp 0 factors,
// the other warps wait on the CTA barrier
__global__ void bench_cta(const double* in, double* out, long long* cycles, int reps) {
  extern __shared__ __align__(16) unsigned char dyn[];
  double (*D)[B + 1] = reinterpret_cast<double (*)[B + 1]>(dyn + 150 * 1024);
  double* col = reinterpret_cast<double*>(dyn + 190 * 1024);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long tot = 0;
  for (int r = 0; r < reps; r++) {
    if (warp == 0) for (int c = 0; c < B; c++) D[lane][c] = in[lane * B + c];
    __syncthreads();
    const long long t0 = clock64();
    if (warp == 0) v3(D, col, lane);
    __syncthreads();
    tot += clock64() - t0;
  }
  if (warp == 0) {
    for (int c = 0; c < B; c++) out[lane * B + c] = (c <= lane) ? D[lane][c] : 0.0;
    if (lane == 0) *cycles = tot / reps;
  }
}
int main() {
  std::vector<double> A(B * B), L(B * B, 0.0);
  for (int i = 0; i < B; i++) for (int j = 0; j <= i; j++) { double v = (i == j) ? 40.0 + i : std::sin(i * 1.3 + j * 0.7); A[i * B + j] = A[j * B + i] = v; }
  for (int j = 0; j < B; j++) { double s = A[j * B + j]; for (int k = 0; k < j; k++) s -= L[j * B + k] * L[j * B + k]; L[j * B + j] = std::sqrt(s);
    for (int i = j + 1; i < B; i++) { double t = A[i * B + j]; for (int k = 0; k < j; k++) t -= L[i * B + k] * L[j * B + k]; L[i * B + j] = t / L[j * B + j]; } }
  double *din, *dout; long long* dc; cudaMalloc(&din, B * B * 8); cudaMalloc(&dout, B * B * 8); cudaMalloc(&dc, 8);
  cudaMemcpy(din, A.data(), B * B * 8, cudaMemcpyHostToDevice);
  auto run = [&](int v) {
    for (int pass = 0; pass < 2; pass++) {
      if (v == 1) bench<1><<<1, 32>>>(din, dout, dc, 20); if (v == 2) bench<2><<<1, 32>>>(din, dout, dc, 20);
      if (v == 3) bench<3><<<1, 32>>>(din, dout, dc, 20); if (v == 4) bench<4><<<1, 32>>>(din, dout, dc, 20);
      if (v == 5) bench<5><<<1, 32>>>(din, dout, dc, 20); if (v == 6) bench<6><<<1, 32>>>(din, dout, dc, 20);
      if (v == 7) bench<7><<<1, 32>>>(din, dout, dc, 20);
      if (v == 13) bench<13><<<1, 32>>>(din, dout, dc, 20); if (v == 17) bench<17><<<1, 32>>>(din, dout, dc, 20);
      cudaDeviceSynchronize();
    }
    std::vector<double> R(B * B); long long c; cudaMemcpy(R.data(), dout, B * B * 8, cudaMemcpyDeviceToHost); cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
    double err = 0; for (int i = 0; i < B * B; i++) err = std::fmax(err, std::fabs(R[i] - L[i]));
    printf("V%d: %lld cycles per 32x32 factorisation (%.2f us @1.965GHz), max |L - L_ref| = %.2e  (%s)\n", v, c, c / 1965.0, err, cudaGetErrorString(cudaGetLastError()));
  };
  for (int v = 1; v <= 7; v++) run(v);
  {
    cudaFuncSetAttribute(bench_cta, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int pass = 0; pass < 2; pass++) { bench_cta<<<1, 256, 200 * 1024>>>(din, dout, dc, 20); cudaDeviceSynchronize(); }
    long long c; cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
    printf("V3 in a 256-thread CTA with 200 KB dynamic smem: %lld cycles (%.2f us) (%s)\n", c, c / 1965.0, cudaGetErrorString(cudaGetLastError()));
    for (int pass = 0; pass < 2; pass++) { bench_cta<<<148, 256, 200 * 1024>>>(din, dout, dc, 20); cudaDeviceSynchronize(); }
    cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
    printf("same, grid 148: %lld cycles (%.2f us) (%s)\n", c, c / 1965.0, cudaGetErrorString(cudaGetLastError()));
  }
  printf("cold instruction cache (V13 = V3, V17 = V7):\n");
  run(13); run(17);
  return 0;
}
