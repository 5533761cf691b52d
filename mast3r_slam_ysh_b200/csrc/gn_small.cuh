// Single-CTA Cholesky + back substitution for small systems (n <= kSmallMaxN, i.e. up to 33 keyframes), entirely in
// shared memory.  At n = 217 (BASELINE configs[2]) the multi-CTA path pays two grid barriers and two L2 round trips per
// 32-column block -- more than the arithmetic; here the packed lower triangle (192 KB) lives in one SM's shared memory
// and the only synchronisation is __syncthreads.  Replaces Eigen::SimplicialLLT on the host (gn_kernels.cu:115-159).
//
// Storage: row r of [H ; b] (row n = rhs) holds columns 0..r, rows padded to an even length so every row starts 16-byte
// aligned: offset P(r) = 2 m (m+1) for r = 2m, 2 (m+1)^2 for r = 2m+1.
#pragma once
#include "gn_internal.cuh"

namespace m3r {

constexpr int kSmallThreads = 256;

__host__ __device__ __forceinline__ int prow(int r) {
  const int m = r >> 1;
  return (r & 1) ? 2 * (m + 1) * (m + 1) : 2 * m * (m + 1);
}

// scratch of the 32x32 diagonal-block kernels below (shared memory)
struct __align__(16) DiagAux {
  double D2[kBlk][kBlk + 2];  // factored diagonal block, zero upper part, identity-padded; even stride for LDS.128
  double col[kBlk];           // pivot column broadcast of the diagonal factorisation
  double inv[kBlk];           // 1 / l_jj
};

struct __align__(16) SmallAux {
  DiagAux d;
  double x[kBlk];             // solution block of the back substitution
  double norm[kSmallThreads / 32];
  int nzrow[kSmallMaxN + 8];  // panel row r (relative to the first row below the block) has a non-zero entry
  int bad;
  long long dbg[64];  // clock64 stamps (diagnostics, see small_mark)
  int dbg_tag[64];
};

__host__ __device__ __forceinline__ size_t small_smem_bytes(int n) {
  return sizeof(double) * (size_t)prow(n + 1) + sizeof(SmallAux);
}

// 1/sqrt(x) on the serial pivot chain: a MUFU seed and two Newton steps in fp64 (error 1.5 e^2 per step: below 1 ulp after
// the second).  x <= 0 or NaN is reported by the caller; a pivot outside the seed's range yields inf/NaN, which the back
// substitution treats as a failed factorisation.
#ifndef M3R_DIAG_V2
#define M3R_DIAG_V2 1
#endif
__device__ __forceinline__ double small_rsqrt(double x) {
#if M3R_DIAG_V2
  // fp64 seed straight from the special-function unit (MUFU.RSQ64H, ~2^-22) instead of F2F -> MUFU.RSQ -> F2F, and the
  // Newton step as y + y (1/2 - (x/2) y^2): three dependent fp64 operations instead of four
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double hx = 0.5 * x;
  y = fma(y, fma(-hx, y * y, 0.5), y);
  y = fma(y, fma(-hx, y * y, 0.5), y);
  return y;
#else
  double y = (double)rsqrtf((float)x);
  const double hx = 0.5 * x;
  y = y * (1.5 - hx * y * y);
  y = y * (1.5 - hx * y * y);
  return y;
#endif
}

__device__ __forceinline__ double small_shfl(double v, int src) {
  int lo = __double2loint(v), hi = __double2hiint(v);
  lo = __shfl_sync(0xffffffffu, lo, src);
  hi = __shfl_sync(0xffffffffu, hi, src);
  return __hiloint2double(hi, lo);
}

// ---- diagonal block: one warp, lane = row held in registers, right-looking, all 32 pivot steps unrolled.  The pivot
//      column travels through shared memory (one STS + broadcast LDS per step); shuffling it costs 3x more
//      (tools/micro/chol32.cu: 3.7 us vs 12 us per block).  A version that keeps only 8 columns in registers and rolls the
//      rest (7 KB of SASS instead of 40 KB) was slower end to end: 122 us vs 95 us per solve at n = 217
//      (tools/variants.sh, r1), so the straight-line code stays.
// myrow: the 32 entries of this lane's row of the block (entries c <= lane are read, and overwritten with L, for
// lane < nb; rows/columns >= nb are identity padding).  d.D2 receives the whole factor, d.inv (and dinv, if given) 1/l_jj.
__device__ __noinline__ bool factor_diag32(double* myrow, DiagAux& d, double* dinv, int nb, int lane) {
  double a[kBlk];
#pragma unroll
  for (int c = 0; c < kBlk; c++) {
    double v = (c == lane) ? 1.0 : 0.0;
    if (lane < nb && c <= lane && c < nb) v = myrow[c];
    a[c] = v;
  }
  bool bad = false;
#if M3R_DIAG_V2
  double djj_next = small_shfl(a[0], 0);
#endif
#pragma unroll
  for (int j = 0; j < kBlk; j++) {
#if M3R_DIAG_V2
    const double djj = djj_next;
#else
    const double djj = small_shfl(a[j], j);
#endif
    if (!(djj > 0.0)) bad = true;  // also catches NaN
    const double inv = small_rsqrt(djj);
    const double l = (lane == j) ? djj * inv : a[j] * inv;
    a[j] = l;
#if M3R_DIAG_V2
    // The next pivot only needs row j + 1's OWN entry of this column: lane j + 1 updates its diagonal from its register
    // and the value goes round by shuffle, so the serial chain is rsqrt -> multiply -> fma -> shuffle and no longer waits
    // for the column's trip through shared memory (which the off-diagonal updates below still take, off the chain).
    // Bitwise the same value as the update below computes for that entry.
    if (j + 1 < kBlk) djj_next = small_shfl(fma(-l, l, a[j + 1]), j + 1);
#endif
    d.col[lane] = l;
    // every lane stores the same value: a lane-0-only store is a divergent region per pivot (BSSY / BRA / BSYNC in the
    // middle of the serial chain), which cost ~100 cycles per pivot (r1)
    d.inv[j] = inv;
    if (dinv != nullptr) dinv[j] = inv;
    __syncwarp();
#pragma unroll
    for (int c = j + 1; c < kBlk; c++) a[c] = fma(-l, d.col[c], a[c]);  // rows above the diagonal: garbage, never read
    __syncwarp();
  }
#pragma unroll
  for (int c = 0; c < kBlk; c++) {
    const double v = (c <= lane) ? a[c] : 0.0;
    if (lane < nb && c <= lane) myrow[c] = v;
    d.D2[lane][c] = v;  // after myrow: in the multi-CTA path myrow aliases D2[lane]
  }
  return bad;
}

// ---- a panel row below the block: x L^T = a in place, row in registers, L broadcast with LDS.128.  p: 16-byte aligned,
//      entries k < nb exist.  Returns whether the row has a non-zero entry (zero rows are left alone).
__device__ __noinline__ int trsm_row32(double* p, const DiagAux& d, int nb) {
  double x[kBlk];
  int nz = 0;
#pragma unroll
  for (int k = 0; k < kBlk; k += 2) {
    double2 v = make_double2(0.0, 0.0);
    if (k + 1 < nb) v = *reinterpret_cast<const double2*>(p + k);
    else if (k < nb) v.x = p[k];
    x[k] = v.x;
    x[k + 1] = v.y;
    nz |= (v.x != 0.0) | (v.y != 0.0);
  }
  if (!nz) return 0;
#pragma unroll
  for (int j = 0; j < kBlk; j++) {
    double s0 = 0.0, s1 = 0.0;
#pragma unroll
    for (int k = 0; k + 1 < j; k += 2) {
      const double2 l = *reinterpret_cast<const double2*>(&d.D2[j][k]);
      s0 = fma(x[k], l.x, s0);
      s1 = fma(x[k + 1], l.y, s1);
    }
    if (j & 1) s0 = fma(x[j - 1], d.D2[j][j - 1], s0);
    x[j] = (x[j] - (s0 + s1)) * d.inv[j];
  }
#pragma unroll
  for (int k = 0; k < kBlk; k += 2) {
    if (k + 1 < nb) *reinterpret_cast<double2*>(p + k) = make_double2(x[k], x[k + 1]);
    else if (k < nb) p[k] = x[k];
  }
  return 1;
}

__device__ __forceinline__ bool small_factor_diag(double* M, SmallAux& ax, double* dinv, int c0, int nb, int lane) {
  return factor_diag32(M + prow(c0 + lane) + c0, ax.d, dinv + c0, nb, lane);
}

__device__ __forceinline__ void small_trsm_row(double* M, SmallAux& ax, int c0, int nb, int row, int rel) {
  ax.nzrow[rel] = trsm_row32(M + prow(row) + c0, ax.d, nb);
}

// ---- trailing update C -= X X^T on the rows t0..n (row n = rhs) / columns t0..n-1.  A warp owns a 16x32 block of C
//      (4x8 threads, 4x4 outputs each): per two k-steps 4 broadcast LDS.128 for the rows and 4 for the columns feed
//      32 DFMA, which keeps the fp64 pipe, not shared memory, the limiter.
// Warp blocks are numbered row block by row block; `worker` of `nworkers` warps takes ids first_id + worker, + nworkers, ...
__device__ __noinline__ void small_update(double* M, const SmallAux& ax, int c0, int n, int first_id, int worker,
                                             int nworkers, int lane) {
  const int t0 = c0 + kBlk;
  const int m = n + 1 - t0;
  if (m <= 0) return;
  const int RB = (m + 15) >> 4;
  const int ty = lane >> 3, tx = lane & 7;
  int w = first_id + worker;
  int base = 0;  // linear id of the first warp block of row block rb
  for (int rb = 0; rb < RB; rb++) {
    const int cnt = (rb >> 1) + 1;  // column blocks cb with 32 cb <= 16 rb + 15
    while (w >= base && w < base + cnt) {
      const int cb = w - base;
      const int r_rel = 16 * rb + 4 * ty, c_rel = 32 * cb + 4 * tx;
      // skip thread tiles that are entirely above the diagonal, outside the matrix, or structurally zero
      bool live = (c_rel <= r_rel + 3) && (r_rel < m) && (c_rel < m - 1);
      int rr[4], cc[4];
      bool any_r = false, any_c = false;
#pragma unroll
      for (int i = 0; i < 4; i++) {
        rr[i] = min(r_rel + i, m - 1);
        cc[i] = min(c_rel + i, m - 1);
        any_r |= ax.nzrow[rr[i]] != 0;
        any_c |= ax.nzrow[cc[i]] != 0;
      }
      live = live && any_r && any_c;
      if (live) {
        const double* ar[4];
        const double* bc[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
          ar[i] = M + prow(t0 + rr[i]) + c0;
          bc[i] = M + prow(t0 + cc[i]) + c0;
        }
        double acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int j = 0; j < 4; j++) acc[i][j] = 0.0;
#pragma unroll 4
        for (int k = 0; k < kBlk; k += 2) {
          double2 a[4], b[4];
#pragma unroll
          for (int i = 0; i < 4; i++) a[i] = *reinterpret_cast<const double2*>(ar[i] + k);
#pragma unroll
          for (int j = 0; j < 4; j++) b[j] = *reinterpret_cast<const double2*>(bc[j] + k);
#pragma unroll
          for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] = fma(a[i].y, b[j].y, fma(a[i].x, b[j].x, acc[i][j]));
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const int r = r_rel + i;
          if (r >= m) continue;
          double* crow = M + prow(t0 + r) + t0;
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const int c = c_rel + j;
            if (c <= r && c < m - 1) crow[c] -= acc[i][j];
          }
        }
      }
      w += nworkers;
    }
    base += cnt;
  }
}

// ---- the whole solve for CTA 0: load [H ; b] from global memory, factor, back-substitute.  Returns with the solution
//      x (L^T x = L^-1 b) in y[0..n) (shared memory, aliases nothing else) and *bad_out set on a failed factorisation.
// diagnostics: clock64 stamps kept in shared memory and copied to the timeline buffer when the solve is over (a mark
// that touches global memory right away slows the code that follows it several-fold and falsifies the picture)
__device__ __forceinline__ void small_mark(SmallAux& ax, int slot, int tag) {
#ifdef M3R_SMALL_DEBUG
  if (threadIdx.x == 0 && slot < 64) {
    ax.dbg_tag[slot] = tag;
    ax.dbg[slot] = clock64();
  }
#endif
}

// Next diagonal block only (rows/columns t0..t0+31 = warp blocks 0 and 1 of small_update), one element per thread: it
// is on the critical path of the look-ahead (the next diagonal factorisation waits for it), so all threads share it.
__device__ __forceinline__ void small_update_diag(double* M, int c0, int n) {
  const int t0 = c0 + kBlk;
  for (int e = threadIdx.x; e < kBlk * kBlk; e += blockDim.x) {
    const int r = e >> 5, c = e & 31;
    if (c > r || t0 + r > n || t0 + c >= n) continue;
    const double* pr = M + prow(t0 + r) + c0;
    const double* pc = M + prow(t0 + c) + c0;
    double s0 = 0.0, s1 = 0.0;
#pragma unroll
    for (int k = 0; k < kBlk; k += 2) {
      const double2 a = *reinterpret_cast<const double2*>(pr + k);
      const double2 b = *reinterpret_cast<const double2*>(pc + k);
      s0 = fma(a.x, b.x, s0);
      s1 = fma(a.y, b.y, s1);
    }
    M[prow(t0 + r) + t0 + c] -= s0 + s1;
  }
}

// sys_packed: [H ; b] in the packed layout of this file (the assembly of small systems writes it that way).
// y, dinv: n (padded to nblk*32) doubles each.  Look-ahead: while warps 1.. apply the trailing update of block column
// kb, warp 0 already factors diagonal block kb+1 (the serial 32-step pivot chain is the critical path of the solve).
__device__ __forceinline__ void small_cholesky_solve(double* M, SmallAux& ax, const double* sys_packed, int n, int nblk,
                                                     double* y, double* dinv, unsigned long long* tl, int tl_cap) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nwarps = blockDim.x >> 5;
  if (tid == 0) ax.bad = 0;
#ifdef M3R_SMALL_DEBUG
  for (int i = tid; i < 64; i += blockDim.x) ax.dbg_tag[i] = 0;
  __syncthreads();
#endif
  small_mark(ax, 0, 1);
  {
    // The first 32 rows first: warp 0 starts on the first diagonal block (whose unrolled code is still cold: ~20 us
    // against ~5 us for the later blocks) while warps 1.. bring in the rest of [H ; b] (4.6 us that used to come before).
    const int total2 = prow(n + 1) >> 1;  // prow is even
    const int head2 = min(total2, prow(min(kBlk, n)) >> 1);
    const double2* src = reinterpret_cast<const double2*>(sys_packed);
    double2* dst = reinterpret_cast<double2*>(M);
    for (int i = tid; i < head2; i += blockDim.x) dst[i] = __ldcg(src + i);
    __syncthreads();
    small_mark(ax, 1, 8);
    if (warp == 0) {
      const bool bad = small_factor_diag(M, ax, dinv, 0, min(kBlk, n), lane);
      if (bad && lane == 0) ax.bad = 1;
    } else {
#pragma unroll 8
      for (int i = head2 + tid - 32; i < total2; i += blockDim.x - 32) dst[i] = __ldcg(src + i);
    }
  }
  __syncthreads();
  for (int kb = 0; kb < nblk; kb++) {
    const int c0 = kb * kBlk;
    const int nb = min(kBlk, n - c0);
    small_mark(ax, 2 + 3 * kb, 1000 + kb);
    const int first = c0 + nb;  // first row below the block; rows first..n
#ifndef M3R_X_NO_TRSM
    for (int row = first + tid; row <= n; row += blockDim.x) small_trsm_row(M, ax, c0, nb, row, row - first);
#endif
    __syncthreads();
    small_mark(ax, 3 + 3 * kb, 2000 + kb);
    if (kb + 1 < nblk) {  // nb == 32 here
#ifdef M3R_SMALL_NO_LOOKAHEAD
      small_update(M, ax, c0, n, 0, warp, nwarps, lane);
      __syncthreads();
      if (warp == 0) {
        const int c1 = c0 + kBlk;
        const bool bad = small_factor_diag(M, ax, dinv, c1, min(kBlk, n - c1), lane);
        if (bad && lane == 0) ax.bad = 1;
      }
      __syncthreads();
      continue;
#endif
#ifndef M3R_X_NO_UPDATE
      small_update_diag(M, c0, n);
#endif
      __syncthreads();
      if (warp == 0) {
#ifndef M3R_X_NO_FACTOR
        const int c1 = c0 + kBlk;
        const bool bad = small_factor_diag(M, ax, dinv, c1, min(kBlk, n - c1), lane);
        if (bad && lane == 0) ax.bad = 1;
#endif
      } else {
#ifndef M3R_X_NO_UPDATE
        small_update(M, ax, c0, n, 2, warp - 1, nwarps - 1, lane);
#endif
      }
      __syncthreads();
    }
    small_mark(ax, 4 + 3 * kb, 3000 + kb);
  }
  // back substitution: y = row n; blocks from the bottom
  for (int c = tid; c < nblk * kBlk; c += blockDim.x) y[c] = (c < n) ? M[prow(n) + c] : 0.0;
  __syncthreads();
#ifdef M3R_X_NO_BACKSOLVE
  if (true) return;
#endif
  for (int kb = nblk - 1; kb >= 0; kb--) {
    const int c0 = kb * kBlk;
    const int nb = min(kBlk, n - c0);
    if (warp == 0) {
      // lane j owns y_j; walk j from the bottom: x_j = y_j / l_jj, then y_i -= l_ji x_j for i < j (row j of L is contiguous)
      double yv = y[c0 + lane];
      const double di = (lane < nb) ? dinv[c0 + lane] : 0.0;
      double lcur = (lane < nb - 1) ? M[prow(c0 + nb - 1) + c0 + lane] : 0.0;  // l_{j,lane} of the step being solved
      for (int j = nb - 1; j >= 0; j--) {
        const double lnext = (j > 0 && lane < j - 1) ? M[prow(c0 + j - 1) + c0 + lane] : 0.0;  // prefetch for step j-1
        const double xj = small_shfl(yv * di, j);
        if (lane == j) yv = xj;
        if (lane < j) yv = fma(-lcur, xj, yv);
        lcur = lnext;
      }
      ax.x[lane] = (lane < nb) ? yv : 0.0;
      y[c0 + lane] = (lane < nb) ? yv : 0.0;
    }
    __syncthreads();
    // y[0:c0] -= L[c0:c0+nb, 0:c0]^T x
    for (int c = tid; c < c0; c += blockDim.x) {
      double acc = 0.0;
#pragma unroll 8
      for (int r = 0; r < nb; r++) acc = fma(M[prow(c0 + r) + c], ax.x[r], acc);
      y[c] -= acc;
    }
    __syncthreads();
  }
#ifdef M3R_SMALL_DEBUG
  small_mark(ax, 63, 7);
  __syncthreads();
  if (tl != nullptr && tid == 0)
    for (int i = 0; i < 64; i++) {
      if (ax.dbg_tag[i] == 0) continue;
      const unsigned long long k = tl[0];
      if ((int)k + 1 < tl_cap) {
        tl[2 * (k + 1)] = 10000 + ax.dbg_tag[i];
        tl[2 * (k + 1) + 1] = (unsigned long long)(ax.dbg[i] - ax.dbg[0]);
        tl[0] = k + 1;
      }
    }
#endif
}

}  // namespace m3r
