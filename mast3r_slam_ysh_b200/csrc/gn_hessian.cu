// Gauss-Newton Hessian assembly for sm_100a: per-edge Sim(3) residuals, Jacobians, IRLS weights and the reduction
// of J^T W J / J^T W r.  Replaces ray_align_kernel / calib_proj_kernel / point_align_kernel of the reference
// (gn_kernels.cu:813-1138, :1231-1543, :455-723) with a different decomposition:
//
//   * The reference accumulates the full 14x14 block per pixel with J_j = Adj(T_i)^-T J_local and J_i = -J_j
//     (gn_kernels.cu:999-1000).  Adj is constant per edge and linear, so we accumulate only
//     S = sum w J_local J_local^T (28 entries) and g = sum w e J_local (7 entries) in the local frame and apply
//     M = A S A^T, v = A g once per edge in fp64 (finalize kernel).  H_ii = H_jj = M, H_ij = H_ji = -M,
//     g_j = v, g_i = -v.
//   * Pose-independent work is hoisted into a packing pre-pass that runs once per solve: points and confidences
//     become one float4 record (one LDG.128 per gather), and idx/valid/Q/C-gates become one 8-byte edge record.
//   * 2-D decomposition: grid = edges x pixel tiles (the reference runs one 256-thread block per edge), fp32
//     warp-shuffle + shared-memory tree per tile, deterministic fp64 second stage across tiles (no atomics).
#include "gn_internal.cuh"

namespace m3r {

// ------------------------------------------------------------------------------------------------ pre-pass
__global__ void __launch_bounds__(256) pack_points_kernel(const float* __restrict__ Xs, const float* __restrict__ Cs,
                                                          float4* __restrict__ xc, long long total) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= total) return;
  xc[p] = make_float4(__ldg(Xs + 3 * p), __ldg(Xs + 3 * p + 1), __ldg(Xs + 3 * p + 2), __ldg(Cs + p));
}

__global__ void __launch_bounds__(256) pack_edges_kernel(const float* __restrict__ Cs, const int64_t* __restrict__ idx,
                                                         const uint8_t* __restrict__ valid, const float* __restrict__ Q,
                                                         const int* __restrict__ ie, const int* __restrict__ je,
                                                         EdgeRec* __restrict__ rec, int num_pts, long long total,
                                                         float C_thresh, float Q_thresh) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= total) return;
  const int e = (int)(g / num_pts);
  const int k = (int)(g - (long long)e * num_pts);
  const int ix = ie[e], jx = je[e];
  EdgeRec r;
  r.idx = -1;
  const float q = __ldg(Q + g);
  r.sq = sqrtf(q);
  if (ix >= 0 && jx >= 0 && valid[g]) {
    const long long ind = idx[g];
    if (ind >= 0 && ind < num_pts) {
      const float ci = __ldg(Cs + (size_t)ix * num_pts + ind);
      const float cj = __ldg(Cs + (size_t)jx * num_pts + k);
      if ((q > Q_thresh) & (ci > C_thresh) & (cj > C_thresh)) r.idx = (int)ind;
    }
  }
  rec[g] = r;
}

// ------------------------------------------------------------------------------------------------ robust weights
template <int ROBUST>
__device__ __forceinline__ float robust_weight(float r) {
  const float a = fabsf(r);
  if (ROBUST == M3R_ROBUST_HUBER) {  // gn_kernels.cu:172-175
    return a < 1.345f ? 1.0f : __fdividef(1.345f, a);
  } else {  // Tukey biweight, nonlinear_optimizer.py:36-42
    const float t = a * (1.0f / 4.6851f);
    const float m = 1.0f - t * t;
    return a < 4.6851f ? m * m : 0.0f;
  }
}

// acc[0..27]: lower triangle of S (row a, col b<=a at a(a+1)/2+b); acc[28..34]: g.
// MASK: bit a set <=> J[a] may be non-zero for this residual row (compile-time sparsity of the local Jacobian).
template <int MASK>
__device__ __forceinline__ void accum_row(float (&acc)[kAcc], const float (&J)[7], float w, float e) {
  float wJ[7];
#pragma unroll
  for (int a = 0; a < 7; a++)
    if ((MASK >> a) & 1) wJ[a] = w * J[a];
  const float we = w * e;
#pragma unroll
  for (int a = 0; a < 7; a++) {
    if (!((MASK >> a) & 1)) continue;
#pragma unroll
    for (int b = 0; b <= a; b++) {
      if (!((MASK >> b) & 1)) continue;
      acc[a * (a + 1) / 2 + b] = fmaf(wJ[a], J[b], acc[a * (a + 1) / 2 + b]);
    }
    acc[28 + a] = fmaf(we, J[a], acc[28 + a]);
  }
}

struct EdgeConst {
  float R[9];  // s_ij * R(q_ij), row-major
  float t[3];
  float fx, fy, cx, cy;
};

// relSim3 (gn_kernels.cu:252-272) folded into a scaled rotation matrix.  The matrix is the linear map of the
// reference's actSO3 formula (X + w*uv + q x uv, uv = 2 q x X), which is what the reference applies even when the
// quaternion has drifted from unit norm (it is never renormalised).
__device__ __forceinline__ void quat_to_mat(const float* q, float* R) {
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  R[0] = 1.0f - 2.0f * (y * y + z * z);
  R[1] = 2.0f * (x * y - w * z);
  R[2] = 2.0f * (x * z + w * y);
  R[3] = 2.0f * (x * y + w * z);
  R[4] = 1.0f - 2.0f * (x * x + z * z);
  R[5] = 2.0f * (y * z - w * x);
  R[6] = 2.0f * (x * z - w * y);
  R[7] = 2.0f * (y * z + w * x);
  R[8] = 1.0f - 2.0f * (x * x + y * y);
}

__device__ void edge_setup(const float* __restrict__ Twc, int ix, int jx, const float* __restrict__ K, EdgeConst* c) {
  const float* Ti = Twc + (size_t)ix * 8;
  const float* Tj = Twc + (size_t)jx * 8;
  const float si_inv = 1.0f / Ti[7];
  const float sij = si_inv * Tj[7];
  const float qi_inv[4] = {-Ti[3], -Ti[4], -Ti[5], Ti[6]};
  const float* qj = Tj + 3;
  float qij[4];
  qij[0] = qi_inv[3] * qj[0] + qi_inv[0] * qj[3] + qi_inv[1] * qj[2] - qi_inv[2] * qj[1];
  qij[1] = qi_inv[3] * qj[1] - qi_inv[0] * qj[2] + qi_inv[1] * qj[3] + qi_inv[2] * qj[0];
  qij[2] = qi_inv[3] * qj[2] + qi_inv[0] * qj[1] - qi_inv[1] * qj[0] + qi_inv[2] * qj[3];
  qij[3] = qi_inv[3] * qj[3] - qi_inv[0] * qj[0] - qi_inv[1] * qj[1] - qi_inv[2] * qj[2];
  float Rij[9], Ri_inv[9];
  quat_to_mat(qij, Rij);
  quat_to_mat(qi_inv, Ri_inv);
  const float d[3] = {Tj[0] - Ti[0], Tj[1] - Ti[1], Tj[2] - Ti[2]};
#pragma unroll
  for (int r = 0; r < 3; r++) {
    c->t[r] = si_inv * (Ri_inv[3 * r] * d[0] + Ri_inv[3 * r + 1] * d[1] + Ri_inv[3 * r + 2] * d[2]);
#pragma unroll
    for (int cc = 0; cc < 3; cc++) c->R[3 * r + cc] = sij * Rij[3 * r + cc];
  }
  if (K != nullptr) {
    c->fx = K[0];
    c->fy = K[4];
    c->cx = K[2];
    c->cy = K[5];
  } else {
    c->fx = c->fy = 1.0f;
    c->cx = c->cy = 0.0f;
  }
}

struct PixParams {
  float sa_inv, sb_inv;  // 1/sigma_a, 1/sigma_b
  int width, height, pixel_border;
  float z_eps;
};

// One pixel of one edge: residuals, weights, local Jacobian rows, accumulation.
template <int KIND, int ROBUST>
__device__ __forceinline__ void pixel_accumulate(float (&acc)[kAcc], const EdgeConst& c, const PixParams& pp,
                                                 const float4 xi, const float4 xj, int ind, float sq) {
  const float Px = fmaf(c.R[0], xj.x, fmaf(c.R[1], xj.y, fmaf(c.R[2], xj.z, c.t[0])));
  const float Py = fmaf(c.R[3], xj.x, fmaf(c.R[4], xj.y, fmaf(c.R[5], xj.z, c.t[1])));
  const float Pz = fmaf(c.R[6], xj.x, fmaf(c.R[7], xj.y, fmaf(c.R[8], xj.z, c.t[2])));
  float J[7];
  if (KIND == M3R_GN_RAYS) {
    const float n2j = fmaf(Px, Px, fmaf(Py, Py, Pz * Pz));
    const float invj = rsqrtf(n2j);
    const float n1j = n2j * invj;
    const float rx = Px * invj, ry = Py * invj, rz = Pz * invj;
    const float n2i = fmaf(xi.x, xi.x, fmaf(xi.y, xi.y, xi.z * xi.z));
    const float invi = rsqrtf(n2i);
    const float n1i = n2i * invi;
    const float e0 = rx - xi.x * invi, e1 = ry - xi.y * invi, e2 = rz - xi.z * invi, e3 = n1j - n1i;
    const float sw_ray = pp.sa_inv * sq, sw_dist = pp.sb_inv * sq;
    const float wc_ray = sw_ray * sw_ray, wc_dist = sw_dist * sw_dist;
    const float w0 = robust_weight<ROBUST>(sw_ray * e0) * wc_ray;
    const float w1 = robust_weight<ROBUST>(sw_ray * e1) * wc_ray;
    const float w2 = robust_weight<ROBUST>(sw_ray * e2) * wc_ray;
    const float w3 = robust_weight<ROBUST>(sw_dist * e3) * wc_dist;
    // d r / d P = (I - r r^T) / |P|   (gn_kernels.cu:982-988)
    const float dxx = invj * (1.0f - rx * rx), dyy = invj * (1.0f - ry * ry), dzz = invj * (1.0f - rz * rz);
    const float dxy = -invj * rx * ry, dxz = -invj * rx * rz, dyz = -invj * ry * rz;
    J[0] = dxx; J[1] = dxy; J[2] = dxz; J[3] = 0.0f; J[4] = rz; J[5] = -ry; J[6] = 0.0f;
    accum_row<0b0110111>(acc, J, w0, e0);
    J[0] = dxy; J[1] = dyy; J[2] = dyz; J[3] = -rz; J[4] = 0.0f; J[5] = rx; J[6] = 0.0f;
    accum_row<0b0101111>(acc, J, w1, e1);
    J[0] = dxz; J[1] = dyz; J[2] = dzz; J[3] = ry; J[4] = -rx; J[5] = 0.0f; J[6] = 0.0f;
    accum_row<0b0011111>(acc, J, w2, e2);
    J[0] = rx; J[1] = ry; J[2] = rz; J[3] = 0.0f; J[4] = 0.0f; J[5] = 0.0f; J[6] = n1j;
    accum_row<0b1000111>(acc, J, w3, e3);
  } else if (KIND == M3R_GN_POINTS) {
    const float e0 = Px - xi.x, e1 = Py - xi.y, e2 = Pz - xi.z;
    const float sw = pp.sa_inv * sq;
    const float wc = sw * sw;
    const float w0 = robust_weight<ROBUST>(sw * e0) * wc;
    const float w1 = robust_weight<ROBUST>(sw * e1) * wc;
    const float w2 = robust_weight<ROBUST>(sw * e2) * wc;
    J[0] = 1.0f; J[1] = 0.0f; J[2] = 0.0f; J[3] = 0.0f; J[4] = Pz; J[5] = -Py; J[6] = Px;
    accum_row<0b1110001>(acc, J, w0, e0);
    J[0] = 0.0f; J[1] = 1.0f; J[2] = 0.0f; J[3] = -Pz; J[4] = 0.0f; J[5] = Px; J[6] = Py;
    accum_row<0b1101010>(acc, J, w1, e1);
    J[0] = 0.0f; J[1] = 0.0f; J[2] = 1.0f; J[3] = Py; J[4] = -Px; J[5] = 0.0f; J[6] = Pz;
    accum_row<0b1011100>(acc, J, w2, e2);
  } else {  // M3R_GN_CALIB, gn_kernels.cu:1360-1479
    const bool valid_z = (Pz > pp.z_eps) && (xi.z > pp.z_eps);
    if (!valid_z) return;  // zero weight in the reference
    const float zj_inv = 1.0f / Pz;
    const float x = Px * zj_inv, y = Py * zj_inv;
    const float u = fmaf(c.fx, x, c.cx), v = fmaf(c.fy, y, c.cy);
    const bool valid_u = (u > (float)pp.pixel_border) && (u < (float)(pp.width - 1 - pp.pixel_border));
    const bool valid_v = (v > (float)pp.pixel_border) && (v < (float)(pp.height - 1 - pp.pixel_border));
    if (!(valid_u && valid_v)) return;
    const int v_t = ind / pp.width;
    const int u_t = ind - v_t * pp.width;
    const float e0 = u - (float)u_t, e1 = v - (float)v_t, e2 = logf(Pz) - logf(xi.z);
    const float sw_pix = pp.sa_inv * sq, sw_dep = pp.sb_inv * sq;
    const float wc_pix = sw_pix * sw_pix, wc_dep = sw_dep * sw_dep;
    const float w0 = robust_weight<ROBUST>(sw_pix * e0) * wc_pix;
    const float w1 = robust_weight<ROBUST>(sw_pix * e1) * wc_pix;
    const float w2 = robust_weight<ROBUST>(sw_dep * e2) * wc_dep;
    J[0] = c.fx * zj_inv; J[1] = 0.0f; J[2] = -c.fx * x * zj_inv; J[3] = -c.fx * x * y;
    J[4] = c.fx * (1.0f + x * x); J[5] = -c.fx * y; J[6] = 0.0f;
    accum_row<0b0111101>(acc, J, w0, e0);
    J[0] = 0.0f; J[1] = c.fy * zj_inv; J[2] = -c.fy * y * zj_inv; J[3] = -c.fy * (1.0f + y * y);
    J[4] = c.fy * x * y; J[5] = c.fy * x; J[6] = 0.0f;
    accum_row<0b0111110>(acc, J, w1, e1);
    J[0] = 0.0f; J[1] = 0.0f; J[2] = zj_inv; J[3] = y; J[4] = -x; J[5] = 0.0f; J[6] = 1.0f;
    accum_row<0b1011100>(acc, J, w2, e2);
  }
}

// ------------------------------------------------------------------------------------------------ Hessian kernel
// grid.x = num_edges * tiles; block = 256 threads; each CTA reduces kTilePix pixels of one edge to kAcc floats.
// ncu on the first version (2048-px tiles, no prefetch) showed it latency-bound, not HBM-bound: long_scoreboard stalls
// on the dependent record -> gather chain at 25 % occupancy, plus 175 shuffles and a serial edge set-up per 8 pixels of
// work (profiles/r1_hessian_v1_ncu_full.txt).  This version therefore
//   * gives every thread kTilePix/256 = 32 pixels (prologue/epilogue amortised 4x),
//   * software-pipelines the loads two stages deep: edge records are fetched two batches ahead and the point records
//     (sequential x_j, gathered x_i) one batch ahead of the arithmetic that consumes them,
//   * reduces 32 of the 35 accumulators with a transposing butterfly (31 shuffles instead of 160).
struct PixBatch {
  float4 xj[2], xi[2];
  int idx[2];
  float sq[2];
};

__device__ __forceinline__ void load_recs(const EdgeRec* __restrict__ er, int kb, int k1, int (&idx)[2], float (&sq)[2]) {
#pragma unroll
  for (int u = 0; u < 2; u++) {
    const int k = kb + u * kHessThreads;
    idx[u] = -1;
    sq[u] = 0.0f;
    if (k < k1) {
      const float2 raw = __ldcs(reinterpret_cast<const float2*>(er + k));  // streamed once: evict-first in L2
      idx[u] = __float_as_int(raw.x);
      sq[u] = raw.y;
    }
  }
}

__device__ __forceinline__ void load_points(const float4* __restrict__ xci, const float4* __restrict__ xcj, int kb,
                                            const int (&idx)[2], const float (&sq)[2], PixBatch& pb) {
#pragma unroll
  for (int u = 0; u < 2; u++) {
    pb.idx[u] = idx[u];
    pb.sq[u] = sq[u];
    if (idx[u] >= 0) {
      pb.xj[u] = __ldg(xcj + kb + u * kHessThreads);
      pb.xi[u] = __ldg(xci + idx[u]);
    }
  }
}

template <int KIND, int ROBUST>
__global__ void __launch_bounds__(kHessThreads, 2) hessian_kernel(const float* __restrict__ Twc,
                                                                  const float* __restrict__ K,
                                                                  const float4* __restrict__ xc,
                                                                  const EdgeRec* __restrict__ rec,
                                                                  const int* __restrict__ ie, const int* __restrict__ je,
                                                                  float* __restrict__ partial,
                                                                  const int* __restrict__ state, int num_pts, int tiles,
                                                                  PixParams pp) {
  if (state[kDone]) return;
  const int e = blockIdx.x / tiles;
  const int tile = blockIdx.x - e * tiles;
  const int ix = ie[e], jx = je[e];
  __shared__ EdgeConst s_c;
  __shared__ float s_red[kHessThreads / 32][kAccStride];

  float acc[kAcc];
#pragma unroll
  for (int i = 0; i < kAcc; i++) acc[i] = 0.0f;

  if (ix >= 0 && jx >= 0) {
    const float4* __restrict__ xci = xc + (size_t)ix * num_pts;
    const float4* __restrict__ xcj = xc + (size_t)jx * num_pts;
    const EdgeRec* __restrict__ er = rec + (size_t)e * num_pts;
    const int k0 = tile * kTilePix;
    const int k1 = min(num_pts, k0 + kTilePix);
    constexpr int kStep = 2 * kHessThreads;
    // stage 0: records of the first two batches are in flight while thread 0 builds the edge constants
    int idx_a[2], idx_b[2];
    float sq_a[2], sq_b[2];
    int kb = k0 + threadIdx.x;
    load_recs(er, kb, k1, idx_a, sq_a);
    load_recs(er, kb + kStep, k1, idx_b, sq_b);
    if (threadIdx.x == 0) edge_setup(Twc, ix, jx, KIND == M3R_GN_CALIB ? K : nullptr, &s_c);
    __syncthreads();
    const EdgeConst c = s_c;
    PixBatch cur, nxt;
    load_points(xci, xcj, kb, idx_a, sq_a, cur);
    for (; kb < k1; kb += kStep) {
      // points of the next batch (its records arrived a stage ago), records two batches ahead
      load_points(xci, xcj, kb + kStep, idx_b, sq_b, nxt);
      load_recs(er, kb + 2 * kStep, k1, idx_b, sq_b);
#pragma unroll
      for (int u = 0; u < 2; u++)
        if (cur.idx[u] >= 0) pixel_accumulate<KIND, ROBUST>(acc, c, pp, cur.xi[u], cur.xj[u], cur.idx[u], cur.sq[u]);
      cur = nxt;
    }
  }

  // ---- block reduction: transposing butterfly for acc[0..31] (lane l ends up with the warp total of acc[l]),
  //      plain xor-tree for the remaining kAcc-32 values, then 8 warps through shared memory
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool upper = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; i++) {
      const float send = upper ? acc[i] : acc[i + s];
      const float keep = upper ? acc[i + s] : acc[i];
      acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  s_red[warp][lane] = acc[0];
#pragma unroll
  for (int i = 32; i < kAcc; i++) {
    float v = acc[i];
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    if (lane == 0) s_red[warp][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < kAcc) {
    float v = 0.0f;
#pragma unroll
    for (int wv = 0; wv < kHessThreads / 32; wv++) v += s_red[wv][threadIdx.x];
    partial[(size_t)blockIdx.x * kAccStride + threadIdx.x] = v;
  }
}

// ------------------------------------------------------------------------------------------------ per-edge finalize
// One warp per edge: sum the tile partials in fp64 (fixed order), M = A S A^T, v = A g with
// A = [[R/s, 0, 0], [[t]x R/s, R, 0], [t^T R/s, 0, 1]] -- the matrix of apply_Sim3_adj_inv(T_i) (gn_kernels.cu:277-297).
__global__ void __launch_bounds__(128) finalize_edges_kernel(const float* __restrict__ Twc,
                                                             const float* __restrict__ partial,
                                                             const int* __restrict__ ie, const int* __restrict__ je,
                                                             double* __restrict__ blocks,
                                                             const int* __restrict__ state, int num_edges, int tiles) {
  if (state[kDone]) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int e = blockIdx.x * 4 + warp;
  __shared__ double sS[4][49], sA[4][49], sB[4][49], sg[4][7];
  if (e >= num_edges) return;
  double* out = blocks + (size_t)e * kBlockDoubles;
  const int ix = ie[e], jx = je[e];
  if (ix < 0 || jx < 0) {
    for (int l = lane; l < kBlockDoubles; l += 32) out[l] = 0.0;
    return;
  }
  double s0 = 0.0, s1 = 0.0;
  const float* p = partial + (size_t)e * tiles * kAccStride;
  for (int t = 0; t < tiles; t++) {
    s0 += (double)p[(size_t)t * kAccStride + lane];
    if (lane < kAcc - 32) s1 += (double)p[(size_t)t * kAccStride + 32 + lane];
  }
  // scatter into symmetric S and g
  {
    int l = lane;  // entries 0..31
    if (l < 28) {
      int a = 0;
      while ((a + 1) * (a + 2) / 2 <= l) a++;
      const int b = l - a * (a + 1) / 2;
      sS[warp][a * 7 + b] = s0;
      sS[warp][b * 7 + a] = s0;
    } else {
      sg[warp][l - 28] = s0;
    }
    if (lane < kAcc - 32) sg[warp][32 + lane - 28] = s1;
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
    double* A = sA[warp];
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
    for (int k = 0; k < 7; k++) v += sA[warp][r * 7 + k] * sS[warp][k * 7 + c];
    sB[warp][l] = v;
  }
  __syncwarp();
  for (int l = lane; l < 49; l += 32) {  // M = B A^T, only lower computed then mirrored for exact symmetry
    const int r = l / 7, c = l % 7;
    if (c <= r) {
      double v = 0.0;
      for (int k = 0; k < 7; k++) v += sB[warp][r * 7 + k] * sA[warp][c * 7 + k];
      out[r * 7 + c] = v;
      out[c * 7 + r] = v;
    }
  }
  if (lane < 7) {
    double v = 0.0;
    for (int k = 0; k < 7; k++) v += sA[warp][lane * 7 + k] * sg[warp][k];
    out[49 + lane] = v;
  }
}

// diagnostic export in the reference's Hs/gs layout (gn_kernels.cu:1097-1137)
__global__ void export_blocks_kernel(const double* __restrict__ blocks, float* __restrict__ Hs, float* __restrict__ gs,
                                     int num_edges) {
  const int e = blockIdx.x;
  const int l = threadIdx.x;
  const double* b = blocks + (size_t)e * kBlockDoubles;
  if (l < 49) {
    const float m = (float)b[l];
    Hs[((size_t)0 * num_edges + e) * 49 + l] = m;
    Hs[((size_t)1 * num_edges + e) * 49 + l] = -m;
    Hs[((size_t)2 * num_edges + e) * 49 + l] = -m;
    Hs[((size_t)3 * num_edges + e) * 49 + l] = m;
  } else if (l < 56) {
    const float v = (float)b[l];
    gs[((size_t)0 * num_edges + e) * 7 + (l - 49)] = -v;
    gs[((size_t)1 * num_edges + e) * 7 + (l - 49)] = v;
  }
}

// ------------------------------------------------------------------------------------------------ launchers
int launch_pack_points(const GnWs& ws, const float* Xs, const float* Cs, cudaStream_t st) {
  const long long total = (long long)ws.L.num_kf * ws.L.num_pts;
  if (total == 0) return M3R_OK;
  pack_points_kernel<<<(unsigned)ceil_div_ll(total, 256), 256, 0, st>>>(Xs, Cs, ws.xc(), total);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int launch_pack_edges(const GnWs& ws, const m3r_gn_params& p, const float* Cs, const int64_t* idx, const uint8_t* valid,
                      const float* Q, cudaStream_t st) {
  const long long total = (long long)ws.L.num_edges * ws.L.num_pts;
  if (total == 0) return M3R_OK;
  pack_edges_kernel<<<(unsigned)ceil_div_ll(total, 256), 256, 0, st>>>(Cs, idx, valid, Q, ws.ie(), ws.je(), ws.rec(),
                                                                       ws.L.num_pts, total, p.C_thresh, p.Q_thresh);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

template <int KIND>
static int launch_hessian_kind(const GnWs& ws, const m3r_gn_params& p, const float* Twc, cudaStream_t st) {
  PixParams pp;
  pp.sa_inv = 1.0f / p.sigma_a;
  pp.sb_inv = (KIND == M3R_GN_POINTS) ? 0.0f : 1.0f / p.sigma_b;
  pp.width = p.width;
  pp.height = p.height;
  pp.pixel_border = p.pixel_border;
  pp.z_eps = p.z_eps;
  const unsigned grid = (unsigned)((long long)ws.L.num_edges * ws.L.tiles);
  if (grid == 0) return M3R_OK;
  if (p.robust == M3R_ROBUST_TUKEY)
    hessian_kernel<KIND, M3R_ROBUST_TUKEY><<<grid, kHessThreads, 0, st>>>(
        Twc, ws.Kdev(), ws.xc(), ws.rec(), ws.ie(), ws.je(), ws.partial(), ws.state(), ws.L.num_pts, ws.L.tiles, pp);
  else
    hessian_kernel<KIND, M3R_ROBUST_HUBER><<<grid, kHessThreads, 0, st>>>(
        Twc, ws.Kdev(), ws.xc(), ws.rec(), ws.ie(), ws.je(), ws.partial(), ws.state(), ws.L.num_pts, ws.L.tiles, pp);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int launch_hessian(const GnWs& ws, const m3r_gn_params& p, const float* Twc, cudaStream_t st) {
  switch (p.kind) {
    case M3R_GN_POINTS: return launch_hessian_kind<M3R_GN_POINTS>(ws, p, Twc, st);
    case M3R_GN_RAYS: return launch_hessian_kind<M3R_GN_RAYS>(ws, p, Twc, st);
    case M3R_GN_CALIB: return launch_hessian_kind<M3R_GN_CALIB>(ws, p, Twc, st);
  }
  set_error("gauss_newton: unknown kind %d", p.kind);
  return M3R_ERR_INVALID_ARGUMENT;
}

int launch_finalize_edges(const GnWs& ws, const float* Twc, cudaStream_t st) {
  if (ws.L.num_edges == 0) return M3R_OK;
  finalize_edges_kernel<<<ceil_div(ws.L.num_edges, 4), 128, 0, st>>>(Twc, ws.partial(), ws.ie(), ws.je(), ws.blocks(),
                                                                      ws.state(), ws.L.num_edges, ws.L.tiles);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int launch_export_blocks(const GnWs& ws, float* Hs, float* gs, cudaStream_t st) {
  if (ws.L.num_edges == 0) return M3R_OK;
  export_blocks_kernel<<<ws.L.num_edges, 64, 0, st>>>(ws.blocks(), Hs, gs, ws.L.num_edges);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

}  // namespace m3r
