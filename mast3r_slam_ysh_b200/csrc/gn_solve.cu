// On-device linear algebra of the Gauss-Newton loop for sm_100a: keyframe index prep, dense fp64 assembly of the
// block system, blocked right-looking Cholesky with the right-hand side carried as an extra row, back substitution,
// Sim(3) retraction and the convergence flag.  Replaces the reference's host path: torch::_unique/searchsorted
// (gn_kernels.cu:161-170), SparseBlock + Eigen::SimplicialLLT in fp64 on the CPU with a D2H/H2D round trip and a
// .item() sync per iteration (:57-159, :1199-1222), and pose_retr_kernel (:415-453).
#include "gn_internal.cuh"

namespace m3r {

// ============================================================================================== index prep
// all[k], k < 2*Ea: concatenation of every rank's ii and jj.  first[k] = 1 iff no earlier equal id.
__global__ void __launch_bounds__(256) uniq_first_kernel(const long long* __restrict__ all, int* __restrict__ first,
                                                         int count) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= count) return;
  const long long v = all[k];
  int f = 1;
  for (int m = 0; m < k; m++)
    if (all[m] == v) {
      f = 0;
      break;
    }
  first[k] = f;
}

// rank of an id among the sorted unique ids == searchsorted(unique, id)  (gn_kernels.cu:166-170)
__global__ void __launch_bounds__(256) edge_rank_kernel(const long long* __restrict__ all, const int* __restrict__ first,
                                                        int count, const int64_t* __restrict__ ii,
                                                        const int64_t* __restrict__ jj, int num_edges, int num_kf,
                                                        int* __restrict__ ie, int* __restrict__ je,
                                                        int* __restrict__ state) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 2 * num_edges) return;
  const int e = t >> 1;
  const long long v = (t & 1) ? (long long)jj[e] : (long long)ii[e];
  int r = 0;
  for (int m = 0; m < count; m++) r += (first[m] && all[m] < v) ? 1 : 0;
  if (r >= num_kf) {  // more distinct keyframe ids than rows of Twc/Xs: the reference would read out of bounds
    r = -1;
    atomicAdd(&state[kBadEdges], 1);
  }
  if (t & 1)
    je[e] = r;
  else
    ie[e] = r;
}

__global__ void concat_ids_kernel(const int64_t* __restrict__ a, const int64_t* __restrict__ b, int n,
                                  long long* __restrict__ out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) {
    out[k] = a[k];
    out[n + k] = b[k];
  }
}

// incidence lists of the LOCAL edges per free pose, in edge order (deterministic assembly order).
// entry = 2*e + role (role 0: pose is the i end, 1: the j end).  Single CTA.
__global__ void __launch_bounds__(1024) incidence_kernel(const int* __restrict__ ie, const int* __restrict__ je,
                                                         int num_edges, int n_free, int* __restrict__ inc_off,
                                                         int* __restrict__ inc_list) {
  extern __shared__ int s_deg[];  // n_free + 1
  for (int a = threadIdx.x; a < n_free; a += blockDim.x) {
    int d = 0;
    for (int e = 0; e < num_edges; e++) {
      const int i = ie[e], j = je[e];
      if (i < 0 || j < 0) continue;
      d += (i - 1 == a) + (j - 1 == a);
    }
    s_deg[a] = d;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int a = 0; a < n_free; a++) {
      const int d = s_deg[a];
      s_deg[a] = run;
      inc_off[a] = run;
      run += d;
    }
    s_deg[n_free] = run;
    inc_off[n_free] = run;
  }
  __syncthreads();
  for (int a = threadIdx.x; a < n_free; a += blockDim.x) {
    int o = s_deg[a];
    for (int e = 0; e < num_edges; e++) {
      const int i = ie[e], j = je[e];
      if (i < 0 || j < 0) continue;
      if (i - 1 == a) inc_list[o++] = 2 * e;
      if (j - 1 == a) inc_list[o++] = 2 * e + 1;
    }
  }
}

__global__ void init_state_kernel(int* state) {
  if (threadIdx.x < kStateInts) state[threadIdx.x] = 0;
}

__global__ void direct_edge_index_kernel(const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int num_edges,
                                         int num_kf, int* __restrict__ ie, int* __restrict__ je,
                                         int* __restrict__ state) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= num_edges) return;
  const long long i = ii[e], j = jj[e];
  const bool ok = i >= 0 && i < num_kf && j >= 0 && j < num_kf;
  if (!ok) atomicAdd(&state[kBadEdges], 1);
  ie[e] = ok ? (int)i : -1;
  je[e] = ok ? (int)j : -1;
}

// ============================================================================================== assembly
// One CTA per free pose a (block row a of H): zero the lower part of the row block and the rhs entries, then add
// the blocks of the incident edges in list order.  Only this CTA writes block row a -> no atomics, fixed order.
//   H[io,io] += M, H[jo,jo] += M, H[io,jo] -= M, H[jo,io] -= M ; b[io] -= v, b[jo] += v   (see gn_hessian.cu)
// System storage: row-major doubles, leading dimension ld, rows 0..n-1 = H (lower triangle valid), row n = rhs.
__global__ void __launch_bounds__(64) assemble_kernel(const double* __restrict__ blocks, const int* __restrict__ ie,
                                                      const int* __restrict__ je, const int* __restrict__ inc_off,
                                                      const int* __restrict__ inc_list, double* __restrict__ sys,
                                                      int* __restrict__ state, int n, int ld) {
  if (state[kDone]) return;
  const int a = blockIdx.x;
  const int t = threadIdx.x;
  if (a == 0 && t == 0) state[kFailNow] = 0;
  // zero
  const int width = 7 * a + 7;
  for (int l = t; l < 7 * width; l += blockDim.x) {
    const int r = l / width, c = l - r * width;
    sys[(size_t)(7 * a + r) * ld + c] = 0.0;
  }
  if (t < 7) sys[(size_t)n * ld + 7 * a + t] = 0.0;
  __syncthreads();
  const int r = t / 7, c = t - r * 7;
  const int beg = inc_off[a], end = inc_off[a + 1];
  double diag = 0.0, rhs = 0.0;
  for (int l = beg; l < end; l++) {
    const int ent = inc_list[l];
    const int e = ent >> 1, role = ent & 1;
    const double* B = blocks + (size_t)e * kBlockDoubles;
    const int other = (role ? ie[e] : je[e]) - 1;  // free index of the other end, -1 if it is the fixed pose
    if (t < 49) {
      const double m = B[t];
      diag += m;
      if (other == a) {
        diag -= m;
      } else if (other >= 0 && other < a) {
        sys[(size_t)(7 * a + r) * ld + 7 * other + c] -= m;
      }
    } else if (t < 56) {
      rhs += role ? B[t] : -B[t];
    }
  }
  if (t < 49) {
    if (c <= r) sys[(size_t)(7 * a + r) * ld + 7 * a + c] = diag;
  } else if (t < 56) {
    sys[(size_t)n * ld + 7 * a + (t - 49)] = rhs;
  }
}

// ============================================================================================== Cholesky
// Blocked right-looking LL^T on the lower triangle, block size 32, rows 0..n (row n = rhs, so forward substitution
// comes for free).  Two launches per block column kb:
//   panel_kernel : every CTA factors the 32x32 diagonal block redundantly in registers (one warp, lane = row), CTA 0
//                  publishes it to lfac[kb]; then each CTA solves X L_kk^T = A for its 64 rows below the block.
//   update_kernel: trailing update C -= X_r X_c^T on 64x64 tiles of the remaining lower triangle.
__device__ __forceinline__ double shfl_d(double v, int src) {
  int lo = __double2loint(v), hi = __double2hiint(v);
  lo = __shfl_sync(0xffffffffu, lo, src);
  hi = __shfl_sync(0xffffffffu, hi, src);
  return __hiloint2double(hi, lo);
}

constexpr int kPanelRows = 64;

__global__ void __launch_bounds__(kPanelRows) panel_kernel(double* __restrict__ sys, double* __restrict__ lfac,
                                                          int* __restrict__ state, int* __restrict__ pflag, int n,
                                                          int ld, int kb) {
  if (state[kDone]) return;
  __shared__ double sL[kBlk][kBlk + 1];
  __shared__ double sInv[kBlk];
  const int c0 = kb * kBlk;
  const int nb = min(kBlk, n - c0);  // valid columns in this block
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp == 0) {
    // lane r holds row r of the diagonal block (lower part); rows/cols >= nb are identity padding
    double a[kBlk];
#pragma unroll
    for (int c = 0; c < kBlk; c++) {
      double v = (c == lane) ? 1.0 : 0.0;
      if (lane < nb && c < nb && c <= lane) v = sys[(size_t)(c0 + lane) * ld + c0 + c];
      a[c] = v;
    }
    bool bad = false;
#pragma unroll
    for (int j = 0; j < kBlk; j++) {
      const double djj = shfl_d(a[j], j);
      if (!(djj > 0.0)) bad = true;  // also catches NaN
      const double inv = rsqrt(djj);  // one MUFU + Newton instead of sqrt + divide on the serial chain
      const double dj = djj * inv;
      const double l = (lane == j) ? dj : a[j] * inv;  // column j of L (valid for lane >= j)
      a[j] = l;
      if (lane == 0) sInv[j] = inv;
#pragma unroll
      for (int c = j + 1; c < kBlk; c++) {
        const double lc = shfl_d(l, c);  // L[c][j]
        if (lane >= c) a[c] -= l * lc;
      }
    }
#pragma unroll
    for (int c = 0; c < kBlk; c++) sL[lane][c] = (c <= lane) ? a[c] : 0.0;
    if (blockIdx.x == 0) {
      double* out = lfac + (size_t)kb * kBlk * kBlk;
#pragma unroll
      for (int c = 0; c < kBlk; c++) out[lane * kBlk + c] = (c <= lane) ? a[c] : 0.0;
      if (bad && lane == 0) state[kFailNow] = 1;
    }
  }
  __syncthreads();
  // rows below the (valid part of the) diagonal block handled by this CTA; row n is the right-hand side
  const int row = c0 + nb + blockIdx.x * kPanelRows + threadIdx.x;
  int nonzero = 0;
  if (row <= n && nb == kBlk) {
    double x[kBlk];
    double* src = sys + (size_t)row * ld + c0;
#pragma unroll
    for (int c = 0; c < kBlk; c++) {
      x[c] = src[c];
      nonzero |= (x[c] != 0.0);
    }
    if (nonzero) {
#pragma unroll
    for (int j = 0; j < kBlk; j++) {
      const double xj = x[j] * sInv[j];
      x[j] = xj;
#pragma unroll
      for (int c = j + 1; c < kBlk; c++) x[c] -= xj * sL[c][j];
    }
#pragma unroll
    for (int c = 0; c < kBlk; c++) src[c] = x[c];
    }
  } else if (row <= n) {  // last, partial block column: only the rhs row can be below it
    nonzero = 1;
    double* src = sys + (size_t)row * ld + c0;
    for (int j = 0; j < nb; j++) {
      const double xj = src[j] * sInv[j];
      src[j] = xj;
      for (int c = j + 1; c < nb; c++) src[c] -= xj * sL[c][j];
    }
  }
  // block-sparsity flag of this 64-row tile of the panel: an all-zero tile stays zero through the TRSM and makes
  // every trailing-update tile that would use it a no-op (0 - 0*x), so update CTAs can exit early
  const int any_nz = __syncthreads_or(nonzero);
  if (threadIdx.x == 0) pflag[blockIdx.x] = any_nz;
}

constexpr int kUpdTile = 64;

// C[r][c] -= sum_k X[r][k] X[c][k], r in row tile, c in col tile, tiles on/below the diagonal of the trailing
// matrix that starts at t0 = (kb+1)*32.  256 threads, 4x4 outputs each.
__global__ void __launch_bounds__(256) update_kernel(double* __restrict__ sys, const int* __restrict__ state,
                                                     const int* __restrict__ pflag, int n, int ld, int kb) {
  if (state[kDone]) return;
  // map linear block id -> (tr, tc) with tc <= tr
  int tr = (int)((sqrtf(8.0f * (float)blockIdx.x + 1.0f) - 1.0f) * 0.5f);
  while ((tr + 1) * (tr + 2) / 2 <= (int)blockIdx.x) tr++;
  while (tr * (tr + 1) / 2 > (int)blockIdx.x) tr--;
  const int tc = blockIdx.x - tr * (tr + 1) / 2;
  if (pflag[tr] == 0 || pflag[tc] == 0) return;  // structurally zero panel tile: nothing to subtract
  const int c0 = kb * kBlk;
  const int t0 = c0 + kBlk;
  const int r_base = t0 + tr * kUpdTile, c_base = t0 + tc * kUpdTile;
  __shared__ double sR[kUpdTile][kBlk + 1];
  __shared__ double sC[kUpdTile][kBlk + 1];
  for (int l = threadIdx.x; l < kUpdTile * kBlk; l += blockDim.x) {
    const int r = l / kBlk, k = l - r * kBlk;
    const int gr = r_base + r, gc = c_base + r;
    sR[r][k] = (gr <= n) ? sys[(size_t)gr * ld + c0 + k] : 0.0;
    sC[r][k] = (gc < n) ? sys[(size_t)gc * ld + c0 + k] : 0.0;
  }
  __syncthreads();
  const int ty = threadIdx.x / 16, tx = threadIdx.x % 16;
  double acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j] = 0.0;
#pragma unroll 8
  for (int k = 0; k < kBlk; k++) {
    double a[4], b[4];
#pragma unroll
    for (int i = 0; i < 4; i++) a[i] = sR[ty + 16 * i][k];
#pragma unroll
    for (int j = 0; j < 4; j++) b[j] = sC[tx + 16 * j][k];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) acc[i][j] = fma(a[i], b[j], acc[i][j]);
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int gr = r_base + ty + 16 * i;
    if (gr > n) continue;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int gc = c_base + tx + 16 * j;
      if (gc < n && gc <= gr) sys[(size_t)gr * ld + gc] -= acc[i][j];
    }
  }
}

// ============================================================================================== Sim(3) retraction
// expSO3 / expSim3 / retrSim3 (gn_kernels.cu:299-413): T <- Exp(xi) o T with xi = (tau, phi, sigma); the
// quaternion is not renormalised.
__device__ __forceinline__ void quat_comp_d(const float* qi, const float* qj, float* out) {
  out[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  out[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  out[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  out[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}

__device__ __forceinline__ void act_so3_d(const float* q, const float* X, float* Y) {
  const float uv0 = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  const float uv1 = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  const float uv2 = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  const float y0 = X[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
  const float y1 = X[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
  const float y2 = X[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
  Y[0] = y0; Y[1] = y1; Y[2] = y2;
}

__device__ void retr_sim3(const float* xi, float* pose) {
  const float kEps = 1e-6f;
  float tau[3] = {xi[0], xi[1], xi[2]};
  const float phi[3] = {xi[3], xi[4], xi[5]};
  const float sigma = xi[6];
  const float scale = expf(sigma);
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float theta = sqrtf(theta_sq);
  float dq[4];
  {
    float imag, real;
    if (theta_sq < kEps) {
      const float theta_p4 = theta_sq * theta_sq;
      imag = 0.5f - (1.0f / 48.0f) * theta_sq + (1.0f / 3840.0f) * theta_p4;
      real = 1.0f - (1.0f / 8.0f) * theta_sq + (1.0f / 384.0f) * theta_p4;
    } else {
      imag = sinf(0.5f * theta) / theta;
      real = cosf(0.5f * theta);
    }
    dq[0] = imag * phi[0]; dq[1] = imag * phi[1]; dq[2] = imag * phi[2]; dq[3] = real;
  }
  float A, B, C;
  if (fabsf(sigma) < kEps) {
    C = 1.0f;
    if (fabsf(theta) < kEps) {
      A = 0.5f;
      B = 1.0f / 6.0f;
    } else {
      A = (1.0f - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - 1.0f) / sigma;
    if (fabsf(theta) < kEps) {
      const float sigma_sq = sigma * sigma;
      A = ((sigma - 1.0f) * scale + 1.0f) / sigma_sq;
      B = (scale * 0.5f * sigma_sq + scale - 1.0f - sigma * scale) / (sigma_sq * sigma);
    } else {
      const float a = scale * sinf(theta);
      const float b = scale * cosf(theta);
      const float c = theta_sq + sigma * sigma;
      A = (a * sigma + (1.0f - b) * theta) / (theta * c);
      B = (C - ((b - 1.0f) * sigma + a * theta) / c) / theta_sq;
    }
  }
  float dt[3] = {C * tau[0], C * tau[1], C * tau[2]};
  float cr[3] = {phi[1] * tau[2] - phi[2] * tau[1], phi[2] * tau[0] - phi[0] * tau[2], phi[0] * tau[1] - phi[1] * tau[0]};
  dt[0] += A * cr[0]; dt[1] += A * cr[1]; dt[2] += A * cr[2];
  float cr2[3] = {phi[1] * cr[2] - phi[2] * cr[1], phi[2] * cr[0] - phi[0] * cr[2], phi[0] * cr[1] - phi[1] * cr[0]};
  dt[0] += B * cr2[0]; dt[1] += B * cr2[1]; dt[2] += B * cr2[2];

  const float t[3] = {pose[0], pose[1], pose[2]};
  const float q[4] = {pose[3], pose[4], pose[5], pose[6]};
  const float s = pose[7];
  float q1[4], t1[3];
  quat_comp_d(dq, q, q1);
  act_so3_d(dq, t, t1);
  pose[0] = t1[0] * scale + dt[0];
  pose[1] = t1[1] * scale + dt[1];
  pose[2] = t1[2] * scale + dt[2];
  pose[3] = q1[0]; pose[4] = q1[1]; pose[5] = q1[2]; pose[6] = q1[3];
  pose[7] = scale * s;
}

__global__ void pose_retr_kernel(float* __restrict__ poses, const float* __restrict__ dx, int num_poses, int num_fix) {
  const int k = num_fix + blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= num_poses) return;
  float xi[7], p[8];
  for (int i = 0; i < 7; i++) xi[i] = dx[(size_t)(k - num_fix) * 7 + i];
  for (int i = 0; i < 8; i++) p[i] = poses[(size_t)k * 8 + i];
  retr_sim3(xi, p);
  for (int i = 0; i < 8; i++) poses[(size_t)k * 8 + i] = p[i];
}

// ============================================================================================== back substitution
// Single CTA: y = row n of sys (= L^-1 b after the factorisation); solve L^T x = y block by block from the bottom,
// dx = -x (gn_kernels.cu:1209), zero on a failed factorisation (:147-150), retraction, ||dx|| < delta test (:1219).
__global__ void __launch_bounds__(1024) backsolve_retract_kernel(const double* __restrict__ sys,
                                                                 const double* __restrict__ lfac,
                                                                 float* __restrict__ Twc, float* __restrict__ dx,
                                                                 int* __restrict__ state, int n, int ld, int nblk,
                                                                 int num_kf, float delta_thresh) {
  if (state[kDone]) return;
  extern __shared__ double s_y[];  // n (+ padding to nblk*32)
  __shared__ double s_x[kBlk];
  __shared__ double s_L[kBlk][kBlk + 1];  // diagonal factor block of the current step
  __shared__ double s_dinv[kBlk];
  __shared__ double s_part[4][256];
  __shared__ double s_norm[32];
  __shared__ int s_bad;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_bad = state[kFailNow];
  for (int c = tid; c < nblk * kBlk; c += blockDim.x) s_y[c] = (c < n) ? sys[(size_t)n * ld + c] : 0.0;
  for (int kb = nblk - 1; kb >= 0; kb--) {
    const int c0 = kb * kBlk;
    {  // stage L_kk (identity-padded, 32x32) in shared memory: the serial loop below must not wait on L2
      const double v = lfac[(size_t)kb * kBlk * kBlk + tid];
      s_L[tid >> 5][tid & 31] = v;
      if ((tid >> 5) == (tid & 31)) s_dinv[tid & 31] = 1.0 / v;
    }
    __syncthreads();
    if (warp == 0) {
      double yv = s_y[c0 + lane];
#pragma unroll
      for (int j = kBlk - 1; j >= 0; j--) {
        const double xj = shfl_d(yv * s_dinv[j], j);  // lane j holds the finished y_j
        if (lane == j) yv = xj;
        if (lane < j) yv -= s_L[j][lane] * xj;
      }
      s_x[lane] = yv;
      s_y[c0 + lane] = yv;
    }
    __syncthreads();
    const int rows = min(kBlk, n - c0);
    // y[0:c0] -= L[c0:c0+rows, 0:c0]^T x : 256 columns x 4 row groups per pass
    for (int cb = 0; cb < c0; cb += 256) {
      const int c = cb + (tid & 255), grp = tid >> 8;
      double acc = 0.0;
      if (c < c0) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
          const int r = grp * 8 + i;
          if (r < rows) acc += sys[(size_t)(c0 + r) * ld + c] * s_x[r];
        }
      }
      s_part[grp][tid & 255] = acc;
      __syncthreads();
      if (tid < 256 && c < c0) s_y[c] -= (s_part[0][tid] + s_part[1][tid]) + (s_part[2][tid] + s_part[3][tid]);
      __syncthreads();
    }
  }
  __syncthreads();
  // non-finite solution counts as a failed factorisation
  int bad_local = 0;
  for (int c = tid; c < n; c += blockDim.x)
    if (!isfinite(s_y[c])) bad_local = 1;
  if (bad_local) atomicOr(&s_bad, 1);
  __syncthreads();
  const bool bad = s_bad != 0;
  double nrm = 0.0;
  for (int c = tid; c < n; c += blockDim.x) {
    const float d = bad ? 0.0f : -(float)s_y[c];
    dx[c] = d;
    nrm += (double)d * (double)d;
  }
  for (int o = 16; o > 0; o >>= 1) {
    const int hi = __shfl_xor_sync(0xffffffffu, __double2hiint(nrm), o);
    const int lo = __shfl_xor_sync(0xffffffffu, __double2loint(nrm), o);
    nrm += __hiloint2double(hi, lo);
  }
  if (lane == 0) s_norm[warp] = nrm;
  __syncthreads();
  // retraction from the shared-memory copy of the solution
  for (int k = tid; k < num_kf - 1; k += blockDim.x) {
    if (!bad) {
      float xi[7], p[8];
      for (int i = 0; i < 7; i++) xi[i] = -(float)s_y[7 * k + i];
      for (int i = 0; i < 8; i++) p[i] = Twc[(size_t)(k + 1) * 8 + i];
      retr_sim3(xi, p);
      for (int i = 0; i < 8; i++) Twc[(size_t)(k + 1) * 8 + i] = p[i];
    }
  }
  if (tid == 0) {
    double tot = 0.0;
    for (int wv = 0; wv < (int)(blockDim.x >> 5); wv++) tot += s_norm[wv];
    state[kIters] += 1;
    if (bad) state[kFailCount] += 1;
    if ((float)sqrt(tot) < delta_thresh) {
      state[kDone] = 1;
      state[kConverged] = 1;
    }
  }
}

// ============================================================================================== launchers
int launch_index_prep(const GnWs& ws, const int64_t* local_ii, const int64_t* local_jj, const int64_t* all_ii,
                      const int64_t* all_jj, int num_all_edges, cudaStream_t st) {
  init_state_kernel<<<1, 32, 0, st>>>(ws.state());
  M3R_LAUNCH_CHECK();
  const int E = ws.L.num_edges;
  const int count = 2 * num_all_edges;
  if (num_all_edges > 0) {
    concat_ids_kernel<<<ceil_div(num_all_edges, 256), 256, 0, st>>>(all_ii, all_jj, num_all_edges, ws.all_ids());
    M3R_LAUNCH_CHECK();
    uniq_first_kernel<<<ceil_div(count, 256), 256, 0, st>>>(ws.all_ids(), ws.uniq_flag(), count);
    M3R_LAUNCH_CHECK();
  }
  if (E > 0) {
    edge_rank_kernel<<<ceil_div(2 * E, 256), 256, 0, st>>>(ws.all_ids(), ws.uniq_flag(), count, local_ii, local_jj, E,
                                                           ws.L.num_kf, ws.ie(), ws.je(), ws.state());
    M3R_LAUNCH_CHECK();
  }
  if (ws.L.n_free > 0) {
    const size_t smem = sizeof(int) * (size_t)(ws.L.n_free + 1);
    if (smem > 200 * 1024) {
      set_error("gauss_newton: too many keyframes (%d) for the incidence pass", ws.L.num_kf);
      return M3R_ERR_UNSUPPORTED;
    }
    if (smem > 48 * 1024)
      M3R_CUDA(cudaFuncSetAttribute(incidence_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    incidence_kernel<<<1, 1024, smem, st>>>(ws.ie(), ws.je(), E, ws.L.n_free, ws.inc_off(), ws.inc_list());
    M3R_LAUNCH_CHECK();
  }
  return M3R_OK;
}

int launch_direct_edge_index(const GnWs& ws, const int64_t* ii_edge, const int64_t* jj_edge, cudaStream_t st) {
  init_state_kernel<<<1, 32, 0, st>>>(ws.state());
  M3R_LAUNCH_CHECK();
  if (ws.L.num_edges == 0) return M3R_OK;
  direct_edge_index_kernel<<<ceil_div(ws.L.num_edges, 256), 256, 0, st>>>(ii_edge, jj_edge, ws.L.num_edges, ws.L.num_kf,
                                                                          ws.ie(), ws.je(), ws.state());
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int launch_assemble(const GnWs& ws, cudaStream_t st) {
  if (ws.L.n_free <= 0) return M3R_OK;
  assemble_kernel<<<ws.L.n_free, 64, 0, st>>>(ws.blocks(), ws.ie(), ws.je(), ws.inc_off(), ws.inc_list(), ws.sys(),
                                              ws.state(), ws.L.n, ws.L.ld);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int launch_cholesky(const GnWs& ws, cudaStream_t st) {
  const int n = ws.L.n, ld = ws.L.ld;
  for (int kb = 0; kb < ws.L.nblk; kb++) {
    const int rb = (kb + 1) * kBlk < n ? (kb + 1) * kBlk : n;
    const int rows_below = n + 1 - rb;  // rows rb .. n (row n = rhs), always >= 1
    const int pgrid = ceil_div(rows_below, kPanelRows);
    panel_kernel<<<pgrid, kPanelRows, 0, st>>>(ws.sys(), ws.lfac(), ws.state(), ws.pflag(), n, ld, kb);
    M3R_LAUNCH_CHECK();
    if ((kb + 1) * kBlk < n) {
      const int T = ceil_div(rows_below, kUpdTile);
      update_kernel<<<T * (T + 1) / 2, 256, 0, st>>>(ws.sys(), ws.state(), ws.pflag(), n, ld, kb);
      M3R_LAUNCH_CHECK();
    }
  }
  return M3R_OK;
}

int launch_backsolve_retract(const GnWs& ws, float* Twc, float delta_thresh, cudaStream_t st) {
  if (ws.L.n_free <= 0) return M3R_OK;
  const size_t smem = sizeof(double) * (size_t)ws.L.nblk * kBlk;
  if (smem > 200 * 1024) {
    set_error("gauss_newton: system of dimension %d exceeds the single-CTA back substitution", ws.L.n);
    return M3R_ERR_UNSUPPORTED;
  }
  if (smem > 40 * 1024)
    M3R_CUDA(cudaFuncSetAttribute(backsolve_retract_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  backsolve_retract_kernel<<<1, 1024, smem, st>>>(ws.sys(), ws.lfac(), Twc, ws.dx(), ws.state(), ws.L.n, ws.L.ld,
                                                  ws.L.nblk, ws.L.num_kf, delta_thresh);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int launch_pose_retr(float* poses, const float* dx, int num_poses, int num_fix, cudaStream_t st) {
  const int cnt = num_poses - num_fix;
  if (cnt <= 0) return M3R_OK;
  pose_retr_kernel<<<ceil_div(cnt, 128), 128, 0, st>>>(poses, dx, num_poses, num_fix);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

}  // namespace m3r
