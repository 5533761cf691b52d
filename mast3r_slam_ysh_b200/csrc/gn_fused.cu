// Fused on-device linear algebra of one Gauss-Newton iteration for sm_100a: ONE cooperative kernel runs
//   finalize (per-edge M = A S A^T in fp64)  ->  zero + assemble the dense lower-triangular [H | b]  ->
//   blocked right-looking Cholesky (rhs carried as row n, so forward substitution is free)  ->
//   back substitution, dx = -x, Sim(3) retraction, convergence flag,
// with grid-wide barriers between the phases instead of kernel boundaries.  It replaces the reference's host path:
// SparseBlock + Eigen::SimplicialLLT in fp64 on the CPU with a D2H/H2D round trip and a .item() sync per iteration
// (gn_kernels.cu:57-159, :1199-1222) and pose_retr_kernel (:415-453).
//
// Why one kernel: at n = 217 (BASELINE configs[2]) the previous version spent 17 launches and ~200 us per iteration on
// kernels whose useful work is a few microseconds each (profiles/r1_launches_v2_summary.txt); the critical path is the
// serial 32-column diagonal-block factorisation, everything else is latency.  A phase mask lets the edge-sharded
// multi-GPU path stop after assembly (all-reduce of [H | b]) and resume with the factorisation.
#include <cooperative_groups.h>

#include "gn_finalize.cuh"
#include "gn_sim3.cuh"
#include "gn_small.cuh"

namespace m3r {

constexpr int kFusedThreads = 256;
constexpr int kPanelRows = kFusedThreads;  // rows of the panel solved by one CTA (one row per thread)
constexpr int kUpdTile = 64;               // trailing-update tile (64x64 outputs per CTA pass)

struct FusedArgs {
  float* Twc;
  const float* partial;
  const long long* eoff;
  const int* ie;
  const int* je;
  double* blocks;
  const int* inc_off;
  const int* inc_list;
  const int* inc_other;
  double* sys;
  double* lfac;
  int* pflag;
  float* dx;
  int* state;
  int n, ld, nblk, num_kf, n_free, num_edges, hess_grid, phases;
  int packed;  // [H ; b] in the packed even-row layout of gn_small.cuh (n <= kSmallMaxN) instead of row-major with ld
  float delta_thresh;
  unsigned long long* timeline;  // diagnostics (m3r_debug_set_timeline): CTA 0 appends (tag, globaltimer ns) pairs
  int timeline_cap;
};

static unsigned long long* g_timeline = nullptr;
static int g_timeline_cap = 0;

__device__ __forceinline__ void tl_mark(const FusedArgs& A, int tag) {
  if (A.timeline != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    const unsigned long long k = A.timeline[0];
    if ((int)k + 1 < A.timeline_cap) {
      A.timeline[2 * (k + 1)] = (unsigned long long)tag;
      A.timeline[2 * (k + 1) + 1] = t;
      A.timeline[0] = k + 1;
    }
  }
}

// ---- grid barrier (sense counter; the last arriver resets the count, so it is reusable across launches) -----------
__device__ __forceinline__ int ld_acquire_i(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_i(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ int ld_relaxed_i(const int* p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Barrier over the first `nb` CTAs of the grid (every participant passes the same nb).  Pollers use relaxed loads with
// a short back-off and fence once after the flip: with acquire loads in the spin loop 148 CTAs serialised on the L2
// line of the counter and one barrier cost ~30 us (measured, r1).
__device__ __forceinline__ void grid_barrier(int* state, int nb) {
  __syncthreads();
  if (threadIdx.x == 0) {
    int* count = state + kBarCount;
    int* gen = state + kBarGen;
    const int g = ld_relaxed_i(gen);
    __threadfence();
    if (atomicAdd(count, 1) == nb - 1) {
      atomicExch(count, 0);
      __threadfence();
      st_release_i(gen, g + 1);
    } else {
      while (ld_relaxed_i(gen) == g) __nanosleep(40);
    }
    __threadfence();
  }
  __syncthreads();
}

__device__ __forceinline__ double shfl_d(double v, int src) {
  int lo = __double2loint(v), hi = __double2hiint(v);
  lo = __shfl_sync(0xffffffffu, lo, src);
  hi = __shfl_sync(0xffffffffu, hi, src);
  return __hiloint2double(hi, lo);
}

// 1/sqrt(x) on the serial pivot chain: fp32 MUFU seed (relative error < 4e-7 including the cast of x) and two Newton
// steps in fp64 (error 1.5 e^2 per step: 2e-13, then below 1 ulp).  x <= 0 or NaN is reported by the caller; a pivot
// outside the fp32 range yields inf/NaN, which the back substitution treats as a failed factorisation.
__device__ __forceinline__ double fast_rsqrt(double x) {
  double y = (double)rsqrtf((float)x);
  const double hx = 0.5 * x;
  y = y * (1.5 - hx * y * y);
  y = y * (1.5 - hx * y * y);
  return y;
}

// ============================================================================================== phase: assemble
// System storage: row-major doubles, leading dimension ld, rows 0..n-1 = H (lower triangle valid), row n = rhs.
//   H[io,io] += M, H[jo,jo] += M, H[io,jo] -= M, H[jo,io] -= M ; b[io] -= v, b[jo] += v   (see gn_hessian.cu)
// Work items, one per 64-thread group, every output written exactly once in a fixed order (no atomics, no
// read-modify-write chains through L2):
//   item a            (a < n_free)        : diagonal block + rhs of free pose a = sum over its incidence list
//   item n_free + l   (l < inc_off[n_free]): off-diagonal block (a, o) of list entry l, o < a; the FIRST entry of a row
//                                            that points at o sums all later entries with the same o
__device__ __forceinline__ size_t sys_at(const FusedArgs& A, int r, int c) {
  return A.packed ? (size_t)(prow(r) + c) : (size_t)r * A.ld + c;
}

__device__ __forceinline__ void assemble_zero(const FusedArgs& A) {
  if (A.packed) {
    const int total2 = prow(A.n + 1) >> 1;
    double2* p = reinterpret_cast<double2*>(A.sys);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total2; i += gridDim.x * blockDim.x) p[i] = make_double2(0.0, 0.0);
    return;
  }
  // zero block row a: columns [0, 7a+7) of rows 7a..7a+6, and the whole rhs row
  const int gwarp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nwarp = gridDim.x * (blockDim.x >> 5);
  const int lane = threadIdx.x & 31;
  for (int r = gwarp; r <= A.n; r += nwarp) {
    const int width = (r == A.n) ? A.n : 7 * (r / 7) + 7;
    double* row = A.sys + (size_t)r * A.ld;
    for (int c = lane; c < width; c += 32) row[c] = 0.0;
  }
}

__device__ __forceinline__ void assemble_items(const FusedArgs& A) {
  const int grp = threadIdx.x >> 6, t = threadIdx.x & 63;
  const int ggrp = blockIdx.x * (blockDim.x >> 6) + grp;
  const int ngrp = gridDim.x * (blockDim.x >> 6);
  const int total_inc = A.inc_off[A.n_free];
  const int r = t / 7, c = t - r * 7;
  constexpr int kOwn = 1 << 30, kMask = kOwn - 1;
  for (int item = ggrp; item < A.n_free + total_inc; item += ngrp) {
    if (item < A.n_free) {
      const int a = item;
      const int beg = A.inc_off[a], end = A.inc_off[a + 1];
      double acc = 0.0;
      if (t < 56) {
#pragma unroll 4
        for (int l = beg; l < end; l++) {  // independent loads: the compiler batches four entries per round trip
          const int ent = A.inc_list[l];
          const int oth = A.inc_other[l];
          const double m = __ldcg(A.blocks + (size_t)(ent >> 1) * kBlockDoubles + t);
          if (t < 49)
            acc += (oth >= 0 && (oth & kMask) == a) ? 0.0 : m;  // a self edge cancels (+M - M)
          else
            acc += (ent & 1) ? m : -m;
        }
      }
      if (t < 49) {
        if (c <= r) A.sys[sys_at(A, 7 * a + r, 7 * a + c)] = acc;
      } else if (t < 56) {
        A.sys[sys_at(A, A.n, 7 * a + (t - 49))] = acc;
      }
    } else {
      const int l = item - A.n_free;
      const int oth = A.inc_other[l];
      if (oth < 0 || !(oth & kOwn)) continue;  // fixed pose, upper triangle, or a later duplicate of an owned block
      const int o = oth & kMask;
      const int ent = A.inc_list[l];
      const int e = ent >> 1;
      const int a = ((ent & 1) ? A.je[e] : A.ie[e]) - 1;
      const int end = A.inc_off[a + 1];
      double acc = 0.0;
      if (t < 49) {
#pragma unroll 4
        for (int l2 = l; l2 < end; l2++) {
          const int oth2 = A.inc_other[l2];
          if (oth2 >= 0 && (oth2 & kMask) == o) acc += __ldcg(A.blocks + (size_t)(A.inc_list[l2] >> 1) * kBlockDoubles + t);
        }
        A.sys[sys_at(A, 7 * a + r, 7 * o + c)] = -acc;
      }
    }
  }
}

// ============================================================================================== phase: Cholesky
// Blocked right-looking LL^T on the lower triangle, block size 32, rows 0..n (row n = rhs).  Per block column kb:
//   panel : CTA i owns rows [c0+nb + 256 i, +256).  The rows and the 32x32 diagonal block are staged in shared memory
//           with coalesced loads; warp 0 factors the diagonal block (factor_diag32, gn_small.cuh), redundantly per CTA,
//           CTA 0 publishes it to lfac[kb]; then every thread solves x L_kk^T = a for its row (trsm_row32) and the
//           rows go back coalesced.  64-row zero flags let the update skip tiles.
//   update: C -= X_r X_c^T on 64x64 tiles of the trailing lower triangle, tiles strided over the CTAs.
// History: the first version broadcast the pivot column with shuffles (12 us per diagonal block, 37 us per panel); a
// rolled shared-memory version took 23 us per panel (latency of un-pipelined LDS chains); see tools/micro/chol32.cu.
struct PanelSmem {
  DiagAux d;                        // diagonal block (staged into d.D2, factored in place), identity-padded beyond nb
  double X[kPanelRows][kBlk + 2];   // this CTA's panel rows (even stride: 16-byte aligned rows for LDS.128)
  int nz[kPanelRows / 64];
};
struct UpdateSmem {
  double R[kUpdTile][kBlk + 1];
  double C[kUpdTile][kBlk + 1];
};

__device__ __forceinline__ void panel_phase(const FusedArgs& A, int kb, PanelSmem& sm) {
  const int n = A.n, ld = A.ld;
  const int c0 = kb * kBlk;
  const int nb = min(kBlk, n - c0);  // valid columns in this block
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int rows_below = n + 1 - (c0 + nb);  // rows c0+nb .. n (row n = rhs), >= 1
  const int ptiles = (rows_below + kPanelRows - 1) / kPanelRows;
  if ((int)blockIdx.x >= ptiles) return;
  const int row0 = c0 + nb + blockIdx.x * kPanelRows;
  __syncthreads();  // previous users of the shared-memory union are done
  // ---- stage: diagonal block (4 rows per warp) and this CTA's rows (32 per warp), lane = column
  for (int i = 0; i < kBlk / (kFusedThreads / 32); i++) {
    const int r = warp * (kBlk / (kFusedThreads / 32)) + i;
    double v = (r == lane) ? 1.0 : 0.0;
    if (r < nb && lane <= r) v = __ldcg(A.sys + (size_t)(c0 + r) * ld + c0 + lane);
    sm.d.D2[r][lane] = v;
  }
#pragma unroll 8
  for (int i = 0; i < 32; i++) {
    const int lr = warp * 32 + i;
    const int row = row0 + lr;
    double v = 0.0;
    if (row <= n && lane < nb) v = __ldcg(A.sys + (size_t)row * ld + c0 + lane);
    sm.X[lr][lane] = v;
  }
  if (threadIdx.x < kPanelRows / 64) sm.nz[threadIdx.x] = 0;
  __syncthreads();
  if (warp == 0) {
    const bool bad = factor_diag32(&sm.d.D2[lane][0], sm.d, nullptr, nb, lane);
    if (blockIdx.x == 0 && bad && lane == 0) A.state[kFailNow] = 1;
  }
  __syncthreads();
  if (blockIdx.x == 0) {  // publish L_kk (zero upper part) for the back substitution
    double* out = A.lfac + (size_t)kb * kBlk * kBlk;
    for (int l = threadIdx.x; l < kBlk * kBlk; l += kFusedThreads) out[l] = sm.d.D2[l >> 5][l & 31];
  }
  // ---- solve x L^T = a for this thread's row; 64-row zero flags: an all-zero tile of the panel stays zero through the
  //      solve and makes every trailing-update tile that would use it a no-op
  if (row0 + (int)threadIdx.x <= n) {
    const int nz = trsm_row32(&sm.X[threadIdx.x][0], sm.d, nb);
    if (nz) atomicOr(&sm.nz[threadIdx.x >> 6], 1);
  }
  __syncthreads();
  if (sm.nz[warp >> 1]) {
#pragma unroll 8
    for (int i = 0; i < 32; i++) {
      const int lr = warp * 32 + i;
      const int row = row0 + lr;
      if (row <= n && lane < nb) A.sys[(size_t)row * ld + c0 + lane] = sm.X[lr][lane];
    }
  }
  if (threadIdx.x < kPanelRows / 64) A.pflag[blockIdx.x * (kPanelRows / 64) + threadIdx.x] = sm.nz[threadIdx.x];
}

// C[r][c] -= sum_k X[r][k] X[c][k] on 64x64 tiles (tr, tc), tc <= tr, of the trailing matrix that starts at
// t0 = c0 + 32; 256 threads, 4x4 outputs each.  A CTA walks its live tiles with a two-deep pipeline: the panel rows of
// the NEXT tile and the old C values of the current one are in flight (registers) while the current tile is multiplied
// out of shared memory, so a tile costs its arithmetic plus one shared-memory round trip instead of three dependent L2
// round trips (load panels -> compute -> read-modify-write C), which made a tile 6 us for 2 us of DFMAs (r1).
struct TileRegs {
  double2 r[4], c[4];
};

__device__ __forceinline__ void tile_decode(int t, int& tr, int& tc) {
  tr = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
  while ((tr + 1) * (tr + 2) / 2 <= t) tr++;
  while (tr * (tr + 1) / 2 > t) tr--;
  tc = t - tr * (tr + 1) / 2;
}

__device__ __forceinline__ void tile_load(const FusedArgs& A, int c0, int tr, int tc, TileRegs& g) {
  const int t0 = c0 + kBlk;
  const int r_base = t0 + tr * kUpdTile, c_base = t0 + tc * kUpdTile;
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int l = threadIdx.x + q * kFusedThreads;
    const int r = l >> 4, k2 = l & 15;
    const int gr = r_base + r, gc = c_base + r;
    g.r[q] = make_double2(0.0, 0.0);
    g.c[q] = make_double2(0.0, 0.0);
    if (gr <= A.n) g.r[q] = __ldcg(reinterpret_cast<const double2*>(A.sys + (size_t)gr * A.ld + c0) + k2);
    if (gc < A.n) g.c[q] = __ldcg(reinterpret_cast<const double2*>(A.sys + (size_t)gc * A.ld + c0) + k2);
  }
}

__device__ __forceinline__ void tile_store(UpdateSmem& sm, const TileRegs& g) {
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int l = threadIdx.x + q * kFusedThreads;
    const int r = l >> 4, k2 = l & 15;
    sm.R[r][2 * k2] = g.r[q].x;
    sm.R[r][2 * k2 + 1] = g.r[q].y;
    sm.C[r][2 * k2] = g.c[q].x;
    sm.C[r][2 * k2 + 1] = g.c[q].y;
  }
}

__device__ __forceinline__ void update_phase(const FusedArgs& A, int kb, int nfac, UpdateSmem& sm) {
  const int n = A.n, ld = A.ld;
  const int c0 = kb * kBlk, t0 = c0 + kBlk;
  const int rows_below = n + 1 - t0;  // rows t0 .. n
  if (rows_below <= 0) return;
  const int T = (rows_below + kUpdTile - 1) / kUpdTile;
  const int ntiles = T * (T + 1) / 2;
  __shared__ int s_pf[128];
  __syncthreads();
  for (int i = threadIdx.x; i < T && i < 128; i += kFusedThreads) s_pf[i] = __ldcg(A.pflag + i);
  __syncthreads();
  auto next_live = [&](int t, int& tr, int& tc) {  // first tile >= t of this CTA whose two panel tiles are non-zero
    for (; t < ntiles; t += nfac) {
      tile_decode(t, tr, tc);
      if (s_pf[tr & 127] != 0 && s_pf[tc & 127] != 0) return t;
    }
    return -1;
  };
  int tr, tc;
  int t = next_live(blockIdx.x, tr, tc);
  if (t < 0) return;
  TileRegs g;
  tile_load(A, c0, tr, tc, g);
  const int ty = threadIdx.x / 16, tx = threadIdx.x % 16;
  while (t >= 0) {
    __syncthreads();  // readers of the previous tile are done with sm
    tile_store(sm, g);
    __syncthreads();
    int ntr = 0, ntc = 0;
    const int tn = next_live(t + nfac, ntr, ntc);
    if (tn >= 0) tile_load(A, c0, ntr, ntc, g);  // in flight during the multiply
    const int r_base = t0 + tr * kUpdTile, c_base = t0 + tc * kUpdTile;
    double cold[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int gr = r_base + ty + 16 * i;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int gc = c_base + tx + 16 * j;
        cold[i][j] = (gr <= n && gc < n && gc <= gr) ? __ldcg(A.sys + (size_t)gr * ld + gc) : 0.0;
      }
    }
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) acc[i][j] = 0.0;
#pragma unroll 8
    for (int k = 0; k < kBlk; k++) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; i++) a[i] = sm.R[ty + 16 * i][k];
#pragma unroll
      for (int j = 0; j < 4; j++) b[j] = sm.C[tx + 16 * j][k];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = fma(a[i], b[j], acc[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int gr = r_base + ty + 16 * i;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int gc = c_base + tx + 16 * j;
        if (gr <= n && gc < n && gc <= gr) A.sys[(size_t)gr * ld + gc] = cold[i][j] - acc[i][j];
      }
    }
    t = tn;
    tr = ntr;
    tc = ntc;
  }
}

// dx = -x (gn_kernels.cu:1209), zero on a failed factorisation (:147-150), retraction (:1212-1215), ||dx|| < delta
// test (:1219).  One CTA; y = solution in shared memory, *s_bad = factorisation failed, s_norm = one slot per warp.
__device__ __forceinline__ void finish_solution(const FusedArgs& A, const double* s_y, int* s_bad, double* s_norm) {
  const int n = A.n;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthr = blockDim.x;
  // non-finite solution counts as a failed factorisation
  int bad_local = 0;
  for (int c = tid; c < n; c += nthr)
    if (!isfinite(s_y[c])) bad_local = 1;
  if (bad_local) atomicOr(s_bad, 1);
  __syncthreads();
  const bool bad = *s_bad != 0;
  double nrm = 0.0;
  for (int c = tid; c < n; c += nthr) {
    const float d = bad ? 0.0f : -(float)s_y[c];
    A.dx[c] = d;
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
  if (!bad) {
    for (int k = tid; k < A.num_kf - 1; k += nthr) {
      float xi[7], p[8];
      for (int i = 0; i < 7; i++) xi[i] = -(float)s_y[7 * k + i];
      for (int i = 0; i < 8; i++) p[i] = A.Twc[(size_t)(k + 1) * 8 + i];
      retr_sim3(xi, p);
      for (int i = 0; i < 8; i++) A.Twc[(size_t)(k + 1) * 8 + i] = p[i];
    }
  }
  if (tid == 0) {
    double tot = 0.0;
    for (int wv = 0; wv < (nthr >> 5); wv++) tot += s_norm[wv];
    A.state[kIters] += 1;
    if (bad) A.state[kFailCount] += 1;
    if ((float)sqrt(tot) < A.delta_thresh) {
      A.state[kDone] = 1;
      A.state[kConverged] = 1;
    }
  }
}

// ============================================================================================== phase: back substitution
// CTA 0 only: y = row n of sys (= L^-1 b after the factorisation); solve L^T x = y block by block from the bottom,
// dx = -x (gn_kernels.cu:1209), zero on a failed factorisation (:147-150), retraction, ||dx|| < delta test (:1219).
struct BackSmem {
  double L[kBlk][kBlk + 1];
  double dinv[kBlk];
  double x[kBlk];
  double norm[kFusedThreads / 32];
  int bad;
};

__device__ __forceinline__ void backsolve_phase(const FusedArgs& A, BackSmem& sm, double* s_y) {
  const int n = A.n, ld = A.ld, nblk = A.nblk;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) sm.bad = __ldcg(A.state + kFailNow);
  for (int c = tid; c < nblk * kBlk; c += kFusedThreads) s_y[c] = (c < n) ? __ldcg(A.sys + (size_t)n * ld + c) : 0.0;
  for (int kb = nblk - 1; kb >= 0; kb--) {
    const int c0 = kb * kBlk;
    __syncthreads();
    for (int l = tid; l < kBlk * kBlk; l += kFusedThreads) {  // stage L_kk (identity-padded) in shared memory
      const double v = __ldcg(A.lfac + (size_t)kb * kBlk * kBlk + l);
      sm.L[l >> 5][l & 31] = v;
      if ((l >> 5) == (l & 31)) sm.dinv[l & 31] = 1.0 / v;
    }
    __syncthreads();
    if (warp == 0) {
      double yv = s_y[c0 + lane];
#pragma unroll
      for (int j = kBlk - 1; j >= 0; j--) {
        const double xj = shfl_d(yv * sm.dinv[j], j);  // lane j holds the finished y_j
        if (lane == j) yv = xj;
        if (lane < j) yv -= sm.L[j][lane] * xj;
      }
      sm.x[lane] = yv;
      s_y[c0 + lane] = yv;
    }
    __syncthreads();
    const int rows = min(kBlk, n - c0);
    // y[0:c0] -= L[c0:c0+rows, 0:c0]^T x : one column per thread, coalesced along the row
    for (int c = tid; c < c0; c += kFusedThreads) {
      double acc = 0.0;
#pragma unroll 8
      for (int r = 0; r < rows; r++) acc += __ldcg(A.sys + (size_t)(c0 + r) * ld + c) * sm.x[r];
      s_y[c] -= acc;
    }
  }
  __syncthreads();
  finish_solution(A, s_y, &sm.bad, sm.norm);
}

// ============================================================================================== the kernel
union FusedSmem {
  FinalizeScratch fin[kFusedThreads / 32];
  PanelSmem panel;
  UpdateSmem upd;
  BackSmem back;
};

__global__ void __launch_bounds__(kFusedThreads, 1) gn_fused_kernel(const FusedArgs A) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  FusedSmem& sm = *reinterpret_cast<FusedSmem*>(s_raw);
  double* s_y = reinterpret_cast<double*>(s_raw + sizeof(FusedSmem));
  // kDone is only written by the last phase of an earlier launch on the same stream: uniform over the grid
  if (__ldcg(A.state + kDone)) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  tl_mark(A, 1);

  if (A.phases & M3R_PH_FINALIZE) {
    const int nwarp = gridDim.x * (kFusedThreads / 32);
    for (int e = blockIdx.x * (kFusedThreads / 32) + warp; e < A.num_edges; e += nwarp)
      finalize_edge(e, lane, sm.fin[warp], A.Twc, A.partial, A.eoff, A.ie, A.je, A.blocks, A.num_edges, A.hess_grid);
  }
  if (A.phases & M3R_PH_ASSEMBLE) {
    if (blockIdx.x == 0 && threadIdx.x == 0) A.state[kFailNow] = 0;
    assemble_zero(A);
    tl_mark(A, 2);
    grid_barrier(A.state, gridDim.x);  // also orders finalize -> assemble
    tl_mark(A, 3);
    assemble_items(A);
    tl_mark(A, 4);
  }
  if (A.phases & M3R_PH_FACTOR) {
    if (A.phases & M3R_PH_ASSEMBLE) grid_barrier(A.state, gridDim.x);
    // only the CTAs that can get work in block column kb take part in its two barriers; the trailing matrix shrinks, so
    // the set is non-increasing and a CTA that drops out never comes back
    tl_mark(A, 5);
    for (int kb = 0; kb < A.nblk; kb++) {
      const int Tk = (A.n + 1 - (kb + 1) * kBlk + kUpdTile - 1) / kUpdTile;
      int nfac = Tk > 0 ? Tk * (Tk + 1) / 2 : 1;
      if (nfac > (int)gridDim.x) nfac = gridDim.x;
      if ((int)blockIdx.x >= nfac) return;
      panel_phase(A, kb, sm.panel);
      tl_mark(A, 1000 + kb);
      grid_barrier(A.state, nfac);
      tl_mark(A, 2000 + kb);
      if ((kb + 1) * kBlk < A.n) {
        update_phase(A, kb, nfac, sm.upd);
        tl_mark(A, 3000 + kb);
        grid_barrier(A.state, nfac);
        tl_mark(A, 4000 + kb);
      }
    }
  }
  if ((A.phases & M3R_PH_BACKSOLVE) && blockIdx.x == 0) {
    backsolve_phase(A, sm.back, s_y);
    tl_mark(A, 6);
  }
}

// ============================================================================================== small systems
// n <= kSmallMaxN: finalize + assembly on all CTAs as above, then CTA 0 factors and solves in shared memory (gn_small.cuh).
__global__ void __launch_bounds__(kSmallThreads, 1) gn_small_kernel(const FusedArgs A) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  if (__ldcg(A.state + kDone)) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  tl_mark(A, 1);
  if (A.phases & M3R_PH_FINALIZE) {
    FinalizeScratch* fin = reinterpret_cast<FinalizeScratch*>(s_raw);
    const int nwarp = gridDim.x * (kSmallThreads / 32);
    for (int e = blockIdx.x * (kSmallThreads / 32) + warp; e < A.num_edges; e += nwarp)
      finalize_edge(e, lane, fin[warp], A.Twc, A.partial, A.eoff, A.ie, A.je, A.blocks, A.num_edges, A.hess_grid);
  }
  if (A.phases & M3R_PH_ASSEMBLE) {
    if (blockIdx.x == 0 && threadIdx.x == 0) A.state[kFailNow] = 0;
    assemble_zero(A);
    tl_mark(A, 2);
    grid_barrier(A.state, gridDim.x);
    tl_mark(A, 3);
    assemble_items(A);
    tl_mark(A, 4);
  }
  if (A.phases & M3R_PH_FACTOR) {
    if (A.phases & M3R_PH_ASSEMBLE) grid_barrier(A.state, gridDim.x);
    if (blockIdx.x != 0) return;
    tl_mark(A, 5);
    double* M = reinterpret_cast<double*>(s_raw);
    SmallAux& ax = *reinterpret_cast<SmallAux*>(M + prow(A.n + 1));
    double* y = reinterpret_cast<double*>(&ax + 1);
    double* dinv = y + A.nblk * kBlk;
    small_cholesky_solve(M, ax, A.sys, A.n, A.nblk, y, dinv, A.timeline, A.timeline_cap);
#ifdef M3R_SMALL_TWICE  // experiment: a second, instruction-cache-warm pass over the same system
    __syncthreads();
    small_cholesky_solve(M, ax, A.sys, A.n, A.nblk, y, dinv, A.timeline, A.timeline_cap);
#endif
    tl_mark(A, 7);
    finish_solution(A, y, &ax.bad, ax.norm);
    tl_mark(A, 6);
  }
}

// ============================================================================================== launcher
int fused_grid(const GnWs& ws) {
  // enough CTAs for the widest phase, never more than one per SM (co-residency is what the grid barrier needs)
  const int n = ws.L.n;
  const int T = ceil_div(n + 1, kUpdTile);
  int want = T * (T + 1) / 2;
  const int fin = ceil_div(ws.L.num_edges, kFusedThreads / 32);
  if (fin > want) want = fin;
  if (want < 4) want = 4;
  const int cap = sm_count();
  return want < cap ? want : cap;
}

int launch_fused(const GnWs& ws, float* Twc, float delta_thresh, int phases, cudaStream_t st) {
  if (ws.L.n_free <= 0) return M3R_OK;
  FusedArgs a;
  a.Twc = Twc;
  a.partial = ws.partial();
  a.eoff = ws.eoff();
  a.ie = ws.ie();
  a.je = ws.je();
  a.blocks = ws.blocks();
  a.inc_off = ws.inc_off();
  a.inc_list = ws.inc_list();
  a.inc_other = ws.inc_other();
  a.sys = ws.sys();
  a.lfac = ws.lfac();
  a.pflag = ws.pflag();
  a.dx = ws.dx();
  a.state = ws.state();
  a.n = ws.L.n;
  a.ld = ws.L.ld;
  a.nblk = ws.L.nblk;
  a.num_kf = ws.L.num_kf;
  a.n_free = ws.L.n_free;
  a.num_edges = ws.L.num_edges;
  a.hess_grid = hessian_grid();
  a.phases = phases;
  a.packed = ws.L.n <= kSmallMaxN ? 1 : 0;
  a.delta_thresh = delta_thresh;
  a.timeline = g_timeline;
  a.timeline_cap = g_timeline_cap;
  void* args[] = {(void*)&a};
  if (ws.L.n <= kSmallMaxN) {
    size_t smem = small_smem_bytes(ws.L.n) + 2 * sizeof(double) * (size_t)ws.L.nblk * kBlk;
    const size_t fin = sizeof(FinalizeScratch) * (kSmallThreads / 32);
    if (smem < fin) smem = fin;
    static size_t attr_small = 0;
    if (smem > 48 * 1024 && smem > attr_small) {
      M3R_CUDA(cudaFuncSetAttribute(gn_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_small = smem;
    }
    int grid = 1;
    if (phases & (M3R_PH_FINALIZE | M3R_PH_ASSEMBLE)) {
      grid = ceil_div(ws.L.num_edges, kSmallThreads / 32);
      const int items = ceil_div(ws.L.n_free + 2 * ws.L.num_edges, kSmallThreads / 64);
      if (items > grid) grid = items;
      if (grid > sm_count()) grid = sm_count();
      if (grid < 1) grid = 1;
    }
    if (grid == 1) {  // no grid barrier inside: a plain launch
      gn_small_kernel<<<1, kSmallThreads, smem, st>>>(a);
      M3R_LAUNCH_CHECK();
      return M3R_OK;
    }
    M3R_CUDA(cudaLaunchCooperativeKernel((const void*)gn_small_kernel, dim3((unsigned)grid), dim3(kSmallThreads), args,
                                         smem, st));
    count_launch();
    return M3R_OK;
  }
  const size_t smem = sizeof(FusedSmem) + sizeof(double) * (size_t)ws.L.nblk * kBlk;
  if (smem > 200 * 1024) {
    set_error("gauss_newton: system of dimension %d exceeds the shared-memory back substitution", ws.L.n);
    return M3R_ERR_UNSUPPORTED;
  }
  static size_t attr_smem = 0;
  if (smem > 48 * 1024 && smem > attr_smem) {
    M3R_CUDA(cudaFuncSetAttribute(gn_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_smem = smem;
  }
  M3R_CUDA(cudaLaunchCooperativeKernel((const void*)gn_fused_kernel, dim3((unsigned)fused_grid(ws)),
                                       dim3(kFusedThreads), args, smem, st));
  count_launch();
  return M3R_OK;
}

}  // namespace m3r

// diagnostics hook (tools/fused_timeline.py): device buffer of 2*cap u64; [0] = number of marks, then (tag, ns) pairs
extern "C" void m3r_debug_set_timeline(unsigned long long* dev_buf, int cap) {
  m3r::g_timeline = dev_buf;
  m3r::g_timeline_cap = cap;
}
