// Per-edge finalize shared by finalize_edges_kernel (diagnostics path) and the fused solver kernel.
#pragma once
#include "gn_internal.cuh"

namespace m3r {

// records per CTA of the persistent Hessian kernel (same formula in the kernel and in finalize_edge)
__host__ __device__ __forceinline__ long long hess_chunk(long long V, int grid) {
  const long long c = (V + grid - 1) / grid;
  return (c + kHessChunkAlign - 1) / kHessChunkAlign * kHessChunkAlign;
}

// scratch of one warp
struct FinalizeScratch {
  double S[49], A[49], B[49], g[7];
};

// One warp per edge: sum the partials of the edge's segments in fp64 (fixed order), M = A S A^T, v = A g with
// A = [[R/s, 0, 0], [[t]x R/s, R, 0], [t^T R/s, 0, 1]] -- the matrix of apply_Sim3_adj_inv(T_i) (gn_kernels.cu:277-297).
__device__ __forceinline__ void finalize_edge(int e, int lane, FinalizeScratch& sc, const float* Twc,
                                              const float* partial, const long long* eoff, const int* ie,
                                              const int* je, double* blocks, int num_edges, int hess_grid) {
  double* out = blocks + (size_t)e * kBlockDoubles;
  const int ix = ie[e], jx = je[e];
  if (ix < 0 || jx < 0) {
    for (int l = lane; l < kBlockDoubles; l += 32) out[l] = 0.0;
    return;
  }
  double s0 = 0.0, s1 = 0.0;
  const long long b0 = eoff[e], b1 = eoff[e + 1];
  if (b1 > b0) {  // segments (e, c) for the CTAs c whose record range intersects [b0, b1): slot e + c
    const long long chunk = hess_chunk(eoff[num_edges], hess_grid);
    const int c_first = (int)(b0 / chunk), c_last = (int)((b1 - 1) / chunk);
    for (int c = c_first; c <= c_last; c++) {
      const float* p = partial + (size_t)(e + c) * kAccStride;
      s0 += (double)__ldcg(p + lane);
      if (lane < kAcc - 32) s1 += (double)__ldcg(p + 32 + lane);
    }
  }
  // scatter into symmetric S and g
  {
    int l = lane;  // entries 0..31
    if (l < 28) {
      int a = 0;
      while ((a + 1) * (a + 2) / 2 <= l) a++;
      const int b = l - a * (a + 1) / 2;
      sc.S[a * 7 + b] = s0;
      sc.S[b * 7 + a] = s0;
    } else {
      sc.g[l - 28] = s0;
    }
    if (lane < kAcc - 32) sc.g[32 + lane - 28] = s1;
  }
  if (lane == 0) {
    const float* T = Twc + (size_t)ix * 8;
    const double t[3] = {T[0], T[1], T[2]};
    const double x = T[3], y = T[4], z = T[5], w = T[6];
    const double s_inv = 1.0 / (double)T[7];
    double R[9];
    R[0] = 1.0 - 2.0 * (y * y + z * z); R[1] = 2.0 * (x * y - w * z); R[2] = 2.0 * (x * z + w * y);
    R[3] = 2.0 * (x * y + w * z); R[4] = 1.0 - 2.0 * (x * x + z * z); R[5] = 2.0 * (y * z - w * x);
    R[6] = 2.0 * (x * z - w * y); R[7] = 2.0 * (y * z + w * x); R[8] = 1.0 - 2.0 * (x * x + y * y);
    double* A = sc.A;
    for (int i = 0; i < 49; i++) A[i] = 0.0;
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) {
        A[r * 7 + c] = s_inv * R[r * 3 + c];
        A[(3 + r) * 7 + 3 + c] = R[r * 3 + c];
      }
    // rows 3..5, cols 0..2: s_inv * [t]x R ; row 6, cols 0..2: s_inv * t^T R
    for (int c = 0; c < 3; c++) {
      const double r0 = R[0 * 3 + c], r1 = R[1 * 3 + c], r2 = R[2 * 3 + c];
      A[3 * 7 + c] = s_inv * (t[1] * r2 - t[2] * r1);
      A[4 * 7 + c] = s_inv * (t[2] * r0 - t[0] * r2);
      A[5 * 7 + c] = s_inv * (t[0] * r1 - t[1] * r0);
      A[6 * 7 + c] = s_inv * (t[0] * r0 + t[1] * r1 + t[2] * r2);
    }
    A[6 * 7 + 6] = 1.0;
  }
  __syncwarp();
  for (int l = lane; l < 49; l += 32) {  // B = A S
    const int r = l / 7, c = l % 7;
    double v = 0.0;
    for (int k = 0; k < 7; k++) v += sc.A[r * 7 + k] * sc.S[k * 7 + c];
    sc.B[l] = v;
  }
  __syncwarp();
  for (int l = lane; l < 49; l += 32) {  // M = B A^T, only lower computed then mirrored for exact symmetry
    const int r = l / 7, c = l % 7;
    if (c <= r) {
      double v = 0.0;
      for (int k = 0; k < 7; k++) v += sc.B[r * 7 + k] * sc.A[c * 7 + k];
      out[r * 7 + c] = v;
      out[c * 7 + r] = v;
    }
  }
  if (lane < 7) {
    double v = 0.0;
    for (int k = 0; k < 7; k++) v += sc.A[lane * 7 + k] * sc.g[k];
    out[49 + lane] = v;
  }
  __syncwarp();
}

}  // namespace m3r
