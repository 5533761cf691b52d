// Microbenchmark 2: small_factor_diag of gn_small.cuh in the solver's launch shape (256 threads, 215 KB dynamic smem,
// warps 1..7 parked at __syncthreads), cold (first call in the kernel) vs warm (later calls), with and without the
// other warps present.
#include <cstdio>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include "../../mast3r_slam_ysh_b200/csrc/gn_small.cuh"
using namespace m3r;
__global__ void __launch_bounds__(256, 1) k(const double* in, long long* cyc, int n, int reps) {
  extern __shared__ __align__(16) unsigned char raw[];
  double* M = reinterpret_cast<double*>(raw);
  SmallAux& ax = *reinterpret_cast<SmallAux*>(M + prow(n + 1));
  double* dinv = reinterpret_cast<double*>(&ax + 1);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int r = 0; r < reps; r++) {
    for (int i = threadIdx.x; i < prow(n + 1); i += blockDim.x) M[i] = in[i];
    __syncthreads();
    long long t0 = clock64();
    if (warp == 0) small_factor_diag(M, ax, dinv, 0, 32, lane);
    __syncthreads();
    if (threadIdx.x == 0) cyc[r] = clock64() - t0;
  }
}
int main() {
  const int n = 217;
  std::vector<double> h(prow(n + 1), 0.0);
  for (int i = 0; i <= n; i++) for (int j = 0; j <= i && j < n; j++) h[prow(i) + j] = (i == j) ? 60.0 + (i % 7) : std::sin(i * 1.3 + j * 0.7) * ((i - j) < 40 ? 1.0 : 0.0);
  double* d; long long* c; cudaMalloc(&d, h.size() * 8); cudaMalloc(&c, 64 * 8);
  cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
  const size_t smem = small_smem_bytes(n) + 2 * 8 * 224;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int threads : {32, 256}) {
    for (int pass = 0; pass < 2; pass++) {
      k<<<1, threads, smem>>>(d, c, n, 6);
      cudaDeviceSynchronize();
      long long hc[6]; cudaMemcpy(hc, c, sizeof(hc), cudaMemcpyDeviceToHost);
      printf("threads=%3d launch %d: cycles per factor call:", threads, pass);
      for (int i = 0; i < 6; i++) printf(" %lld", hc[i]);
      printf("  (%s)\n", cudaGetErrorString(cudaGetLastError()));
    }
  }
  return 0;
}
