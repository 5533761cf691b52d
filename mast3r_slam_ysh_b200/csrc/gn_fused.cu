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
#include "nd_plan.cuh"
#include "tma.cuh"

namespace m3r {

constexpr int kFusedThreads = 256;
constexpr int kPanW = 64;                  // panel width: two 32-column steps inside one phase
constexpr int kPanelRows = 192;            // rows of the panel solved by one CTA (six row warps, one row per thread)
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
  int* pmask;   // [panel][64-row group] non-zero flags of the factor (for the back substitution), stride pstride
  int pstride;
  const int* sky;  // envelope plan (SkyPlan), nullptr when the envelope solver cannot run (shared memory, debug switch)
  const int* nd;   // segmented schedule of the envelope solver (NdPlan, nd_plan.cuh), nullptr when switched off
  double* ndbuf;   // its scratch: Schur complements and right-hand sides of the interiors, the solution (NdBuf)
  int sy_off;      // gn_fused_kernel: byte offset of the solution vector behind the phase union in dynamic shared memory
  float* dx;
  int* state;
  int n, ld, nblk, num_kf, n_free, num_edges, hess_grid, phases;
  int packed;  // [H ; b] in the packed even-row layout of gn_small.cuh (n <= kSmallMaxN) instead of row-major with ld
  float delta_thresh;
  unsigned long long* timeline;  // diagnostics (m3r_debug_set_timeline): CTA 0 appends (tag, globaltimer ns) pairs
  int timeline_cap;
  PeerInfo peer;  // world > 1: sum [H | b] over the ranks inside the kernel (peer_allreduce_phase)
};

extern int g_solver_mode;  // gn_solve.cu
static unsigned long long* g_timeline = nullptr;
static int g_timeline_cap = 0;

__device__ __forceinline__ void tl_mark(const FusedArgs& A, int tag) {
  if (A.timeline != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    const unsigned long long k = atomicAdd(A.timeline, 1ULL);
    if ((int)k + 1 < A.timeline_cap) {
      A.timeline[2 * (k + 1)] = (unsigned long long)tag;
      A.timeline[2 * (k + 1) + 1] = t;
    }
  }
}

// same, from any one thread of CTA 0 (slots claimed atomically: marks of different warps may interleave)
__device__ __forceinline__ void tl_mark_any(const FusedArgs& A, int tag) {
  if (A.timeline != nullptr && blockIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    const unsigned long long k = atomicAdd(A.timeline, 1ULL);
    if ((int)k + 1 < A.timeline_cap) {
      A.timeline[2 * (k + 1)] = (unsigned long long)tag;
      A.timeline[2 * (k + 1) + 1] = t;
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

// ============================================================================================== phase: peer all-reduce
// Multi-GPU: every rank has assembled the [H | b] of ITS edges into its own buffer in peer-mapped memory.  Between
// assembly and factorisation the fused kernel sums the buffers itself over NVLink:
//   barrier (all local CTAs + all ranks: assembly done, visible system-wide)
//   -> rank r owns every world-th row (small systems: every world-th 256-chunk of the packed triangle); its CTAs read
//      that slice from all ranks (peer loads), add in rank order and write the sum into ALL ranks' buffers (peer stores)
//   -> barrier (all sums landed) -> factorisation.
// Only the lower triangle + rhs row is touched.  Each element is summed by exactly one rank and then replicated, so
// the replicas stay bit-identical, as with the NCCL all-reduce this replaces (one launch + one collective call less per
// iteration).  A rank that never arrives (crashed peer) is reported through state[kPeerTimeout] after ~4 s instead of
// hanging the device.
__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Grid barrier over all CTAs of this launch; the CTA that arrives last also meets the other ranks at value `v`.
// Warp 0 of every CTA takes part: lane 0 does the local arrival, and in the last CTA LANE p talks to PEER p -- the
// signals go out together and the flags are polled together (one fence before the relaxed signal stores, one after the
// relaxed polls).  The first version signalled and polled the peers one after the other from a single thread with
// release stores / acquire loads: ~27 us per barrier at 8 ranks, 120 us for the whole all-reduce phase against 45 us
// at 2 ranks (tools/multi_timeline.py).
__device__ __forceinline__ void grid_peer_barrier(const FusedArgs& A, unsigned long long v) {
  __syncthreads();
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    int* count = A.state + kBarCount;
    int* gen = A.state + kBarGen;
    int g = 0, last = 0;
    if (lane == 0) {
      g = ld_relaxed_i(gen);
      __threadfence_system();  // this CTA's writes, peer stores included, are visible system-wide before it arrives
      last = atomicAdd(count, 1) == (int)gridDim.x - 1;
      if (last) atomicExch(count, 0);
    }
    last = __shfl_sync(0xffffffffu, last, 0);
    if (last) {
      const PeerInfo& P = A.peer;
      __syncwarp();
      if (lane < P.world && lane != P.rank) {
        __threadfence_system();  // cumulative: everything the arrivals ordered before lane 0's atomic comes before the signal
        st_relaxed_sys_u64(P.flags[lane] + P.rank, v);
        long long spins = 0;
        while (ld_relaxed_sys_u64(P.flags[P.rank] + lane) < v) {
          __nanosleep(64);
          if (++spins > (1LL << 25)) {  // ~4 s: give up instead of hanging the GPU
            A.state[kPeerTimeout] = 1;
            break;
          }
        }
        __threadfence_system();  // the peer's writes before its signal are visible from here on
      }
      __syncwarp();
      if (lane == 0) {
        __threadfence_system();
        st_release_i(gen, g + 1);
      }
    } else if (lane == 0) {
      while (ld_relaxed_i(gen) == g) __nanosleep(40);
    }
    if (lane == 0) __threadfence_system();
  }
  __syncthreads();
}

__device__ __forceinline__ void peer_sum_store(const FusedArgs& A, size_t off2) {  // off2: index in double2 units
  const PeerInfo& P = A.peer;
  double2 acc = make_double2(0.0, 0.0);
#pragma unroll
  for (int p = 0; p < kMaxPeers; p++) {
    if (p >= P.world) break;
    double2 v;  // system-scope accesses: peer memory is coherent at the home GPU's L2, never cached locally
    asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];"
                 : "=d"(v.x), "=d"(v.y)
                 : "l"(reinterpret_cast<const double2*>(P.sys[p]) + off2)
                 : "memory");
    acc.x += v.x;
    acc.y += v.y;
  }
#pragma unroll
  for (int p = 0; p < kMaxPeers; p++) {
    if (p >= P.world) break;
    asm volatile("st.relaxed.sys.global.v2.f64 [%0], {%1, %2};" ::"l"(reinterpret_cast<double2*>(P.sys[p]) + off2),
                 "d"(acc.x), "d"(acc.y)
                 : "memory");
  }
}

__device__ __forceinline__ int row_first_col(const FusedArgs& A, int sky, int r);

template <int THREADS>
__device__ __forceinline__ void peer_allreduce_phase(const FusedArgs& A, int sky) {
  const PeerInfo& P = A.peer;
  grid_peer_barrier(A, 2 * P.epoch + 1);
  if (A.packed) {  // packed lower triangle + rhs row: one linear range of prow(n + 1) doubles (even by construction)
    const size_t n2 = (size_t)prow(A.n + 1) / 2;
    const size_t nchunks = (n2 + THREADS - 1) / THREADS;
    for (size_t c = (size_t)P.rank + (size_t)P.world * blockIdx.x; c < nchunks; c += (size_t)P.world * gridDim.x) {
      const size_t i = c * THREADS + threadIdx.x;
      if (i < n2) peer_sum_store(A, i);
    }
  } else {  // row-major with leading dimension ld (even): row r holds columns row_first_col(r)..min(r, n-1)
    // Work items = (row of this rank, chunk of THREADS double2), chunk-major: a thread makes ONE remote round trip per
    // item and a CTA gets at most two or three non-empty items.  (One CTA per row, looping over the row, gave the threads
    // of the long rows at the bottom of a ring's envelope four or five round trips in a row while most of the grid idled.)
    const int rows = (A.n - P.rank) / P.world + 1;              // rows P.rank, P.rank + world, ... <= n
    const int chunks = ((A.n + 1) / 2 + THREADS - 1) / THREADS;  // of the longest row
    for (int item = blockIdx.x; item < rows * chunks; item += gridDim.x) {
      const int chunk = item / rows;
      const int r = P.rank + P.world * (item - chunk * rows);
      const int len2 = (min(r + 1, A.n) + 1) / 2;
      const int i = row_first_col(A, sky, r) / 2 + chunk * THREADS + threadIdx.x;
      if (i < len2) peer_sum_store(A, (size_t)r * A.ld / 2 + i);
    }
  }
  grid_peer_barrier(A, 2 * P.epoch + 2);
}

// A peer never arrived: this rank holds a PARTIAL sum.  Nothing may be factored or retracted from it -- the solve stops
// here (kDone, dx = 0, poses untouched) and m3r_gn_end reports bit 30 of info[3], which the host side turns into an
// error.  The flag is written before the barrier releases, so every CTA of the launch reads the same value.
__device__ __forceinline__ bool peer_timed_out(const FusedArgs& A) {
  if (__ldcg(A.state + kPeerTimeout) == 0) return false;
  if (blockIdx.x == 0) {
    for (int i = threadIdx.x; i < A.n; i += blockDim.x) A.dx[i] = 0.0f;
    if (threadIdx.x == 0) {
      A.state[kFailCount] += 1;
      A.state[kDone] = 1;
    }
  }
  return true;
}

// 16-byte global -> shared copy through L2; bytes beyond src_bytes (0..16) are zero-filled
__device__ __forceinline__ void cp_async16_zfill(void* smem, const void* gmem, int src_bytes) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sa), "l"(gmem), "r"(src_bytes) : "memory");
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

// First column of row r that the solver will ever read: with the envelope solver nothing left of the row block's
// envelope is touched (not by the assembly, the joins or the back substitution), so it is neither zeroed nor summed
// over the ranks (12.7 MB -> 1.8 MB per iteration at n = 1785 with a window of 8).
__device__ __forceinline__ int row_first_col(const FusedArgs& A, int sky, int r) {
  return (sky > 0 && r < A.n) ? kBlk * __ldg(A.sky + SkyPlan::fb(A.n_free) + r / kBlk) : 0;
}

__device__ __forceinline__ void assemble_zero(const FusedArgs& A, int sky) {
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
    for (int c = row_first_col(A, sky, r) + lane; c < width; c += 32) row[c] = 0.0;
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
// Blocked right-looking LL^T on the lower triangle, rows 0..n (row n = rhs).  The panel is 64 columns wide and is
// factored as two 32-column steps INSIDE one phase, so the grid pays one panel, one update and two barriers per 64
// columns (a 32-column panel cost 28 us per block column at n = 1785, most of it fixed latencies: staging, barriers,
// the update's L2 round trips; r1).  Per panel kp (columns c0 = 64 kp, w <= 64 of them valid):
//   panel : CTA i owns rows [c0+w + 192 i, +192): six row warps, one row per thread, rows staged in shared memory.
//           Warp 6 owns the 64x64 diagonal block [[D11, 0], [D21, D22]], redundantly per CTA:
//             factor D11 | L21 = D21 L11^-T | D22 -= L21 L21^T, factor D22
//           while the row warps run   X1 = X1 L11^-T | X2 -= X1 L21^T | X2 = X2 L22^-T   with three CTA barriers in
//           between (factor_diag32 / trsm_row32, gn_small.cuh).  CTA 0 publishes L11, L22 to lfac and L21 to sys for the
//           back substitution.  64-row zero flags let the update skip tiles.
//   update: C -= X_r X_c^T (K = 64) on 64x64 tiles of the trailing lower triangle, tiles strided over the CTAs.
// History: the first version broadcast the pivot column with shuffles (12 us per diagonal block, 37 us per panel); a
// rolled shared-memory version took 23 us per panel (latency of un-pipelined LDS chains); see tools/micro/chol32.cu.
constexpr int kDiagWarp = kPanelRows / 32;  // warp 6 (warp 7 only helps staging)
struct PanelSmem {
  DiagAux d1, d2;                    // D11 / D22, staged then factored in place, identity-padded
  double L21[kBlk][kBlk + 2];        // D21 -> L21
  double X[kPanelRows][kPanW + 2];   // this CTA's panel rows (stride = 16 mod 128 bytes: conflict-free LDS.128 per row)
  int nz[kPanelRows / 64];
};
struct UpdateSmem {
  double R[kUpdTile][kPanW + 1];
  double C[kUpdTile][kPanW + 1];
};

// dst[c] -= sum_k a[k] * L[c][k] for c = cstart, cstart + cstep, ... < climit; a = this thread's 32 values (registers),
// L rows broadcast from shared memory, dst = this thread's row in shared memory
__device__ __forceinline__ void row_minus_a_Lt(const double (&a)[kBlk], const double (*L)[kBlk + 2], double* dst,
                                               int climit, int cstart = 0, int cstep = 1) {
#pragma unroll 2
  for (int c = cstart; c < climit; c += cstep) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
    for (int k = 0; k < kBlk; k += 4) {
      const double2 l0 = *reinterpret_cast<const double2*>(&L[c][k]);
      const double2 l1 = *reinterpret_cast<const double2*>(&L[c][k + 2]);
      s0 = fma(a[k], l0.x, s0);
      s1 = fma(a[k + 1], l0.y, s1);
      s2 = fma(a[k + 2], l1.x, s2);
      s3 = fma(a[k + 3], l1.y, s3);
    }
    dst[c] -= (s0 + s1) + (s2 + s3);
  }
}

#ifdef M3R_PANEL_MARKS
#define PMARK(t) tl_mark(A, (t) + (kp == 3 ? 0 : 50))
#else
#define PMARK(t)
#endif
__device__ __forceinline__ void panel_phase(const FusedArgs& A, int kp, PanelSmem& sm) {
  const int n = A.n, ld = A.ld;
  const int c0 = kp * kPanW;
  const int w = min(kPanW, n - c0);  // valid columns of this panel
  const int nb1 = min(kBlk, w), nb2 = w - nb1;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rows_below = n + 1 - (c0 + w);  // rows c0+w .. n (row n = rhs), >= 1
  const int ptiles = (rows_below + kPanelRows - 1) / kPanelRows;
  if ((int)blockIdx.x >= ptiles) return;
  const int row0 = c0 + w + blockIdx.x * kPanelRows;
  __syncthreads();  // previous users of the shared-memory union are done
  // ---- stage (lane = column): warps 6, 7 the diagonal block, the row warps their 32 rows each
  if (warp >= kDiagWarp) {
    const int half = warp - kDiagWarp;
#pragma unroll
    for (int i = 0; i < kBlk / 2; i++) {
      const int r = half * (kBlk / 2) + i;
      double v = (r == lane) ? 1.0 : 0.0;
      if (r < nb1 && lane <= r) v = __ldcg(A.sys + (size_t)(c0 + r) * ld + c0 + lane);
      sm.d1.D2[r][lane] = v;
      v = 0.0;
      if (r < nb2 && lane < nb1) v = __ldcg(A.sys + (size_t)(c0 + kBlk + r) * ld + c0 + lane);
      sm.L21[r][lane] = v;
      v = (r == lane) ? 1.0 : 0.0;
      if (r < nb2 && lane <= r) v = __ldcg(A.sys + (size_t)(c0 + kBlk + r) * ld + c0 + kBlk + lane);
      sm.d2.D2[r][lane] = v;
    }
  } else {
    // 32 rows x 64 columns per warp, a 16-byte cp.async per lane and row, all in flight at once (staging with
    // register loads, 16 in flight, took 8.8 us of a 33 us panel at n = 1785; r1).  Bytes beyond the valid columns
    // / rows are zero-filled by the copy itself.
#pragma unroll
    for (int i = 0; i < 32; i++) {
      const int lr = warp * 32 + i;
      const int row = row0 + lr;
      const int col = 2 * lane;
      int bytes = row <= n ? (w - col) * 8 : 0;
      bytes = bytes < 0 ? 0 : (bytes > 16 ? 16 : bytes);
      const double* src = bytes > 0 ? A.sys + (size_t)row * ld + c0 + col : A.sys;
      cp_async16_zfill(&sm.X[lr][col], src, bytes);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  if (tid < kPanelRows / 64) sm.nz[tid] = 0;
  __syncthreads();
  PMARK(5100);
  // ---- step 1: L11
  if (warp == kDiagWarp) {
    const bool bad = factor_diag32(&sm.d1.D2[lane][0], sm.d1, nullptr, nb1, lane);
    if (blockIdx.x == 0 && bad && lane == 0) A.state[kFailNow] = 1;
  }
  __syncthreads();
  PMARK(5200);
  // ---- step 2: L21 = D21 L11^-T (warp 6)  |  X1 = X1 L11^-T (row warps)
  const bool has_row = warp < kDiagWarp && row0 + tid <= n;
  int nz1 = 0;
  if (warp == kDiagWarp) {
    trsm_row32(&sm.L21[lane][0], sm.d1, nb1);
  } else if (has_row) {
    nz1 = trsm_row32(&sm.X[tid][0], sm.d1, nb1);
  }
  __syncthreads();
  PMARK(5300);
  // ---- step 3: D22 -= L21 L21^T, factor (warp 6)  |  X2 -= X1 L21^T (row warps)
  if (warp >= kDiagWarp) {
    if (nb2 > 0) {  // warps 6 and 7 split the columns of the 32x32 update (even / odd), then meet on barrier 1
      double a[kBlk];
#pragma unroll
      for (int k = 0; k < kBlk; k += 2) {
        const double2 v = *reinterpret_cast<const double2*>(&sm.L21[lane][k]);
        a[k] = v.x;
        a[k + 1] = v.y;
      }
      row_minus_a_Lt(a, sm.L21, &sm.d2.D2[lane][0], lane < nb2 ? lane + 1 : 0, warp - kDiagWarp, 2);
      asm volatile("bar.sync 1, 64;" ::: "memory");
    }
    if (warp == kDiagWarp) {
      const bool bad = factor_diag32(&sm.d2.D2[lane][0], sm.d2, nullptr, nb2, lane);
      if (blockIdx.x == 0 && bad && lane == 0) A.state[kFailNow] = 1;
    }
  } else if (has_row && nz1 && nb2 > 0) {
    double a[kBlk];
#pragma unroll
    for (int k = 0; k < kBlk; k += 2) {
      const double2 v = *reinterpret_cast<const double2*>(&sm.X[tid][k]);
      a[k] = v.x;
      a[k + 1] = v.y;
    }
    row_minus_a_Lt(a, sm.L21, &sm.X[tid][kBlk], nb2);
  }
  __syncthreads();
  PMARK(5400);
  // ---- step 4: X2 = X2 L22^-T; 64-row zero flags: an all-zero tile of the panel stays zero through the solves and
  //      makes every trailing-update tile that would use it a no-op
  if (has_row) {
    int nz = nz1;
    if (nb2 > 0) nz |= trsm_row32(&sm.X[tid][kBlk], sm.d2, nb2);
    if (nz) atomicOr(&sm.nz[tid >> 6], 1);
  }
  __syncthreads();
  PMARK(5500);
  if (warp < kDiagWarp) {
    if (sm.nz[warp >> 1]) {
#pragma unroll 8
      for (int i = 0; i < 32; i++) {
        const int lr = warp * 32 + i;
        const int row = row0 + lr;
        if (row <= n) {
          if (lane < w) A.sys[(size_t)row * ld + c0 + lane] = sm.X[lr][lane];
          if (kBlk + lane < w) A.sys[(size_t)row * ld + c0 + kBlk + lane] = sm.X[lr][kBlk + lane];
        }
      }
    }
  } else if (blockIdx.x == 0) {  // publish L11, L22 (zero upper parts) and L21 for the back substitution
    const int t = tid - kDiagWarp * 32;  // 0..63
    double* out1 = A.lfac + (size_t)(2 * kp) * kBlk * kBlk;
    for (int l = t; l < kBlk * kBlk; l += 64) out1[l] = sm.d1.D2[l >> 5][l & 31];
    if (nb2 > 0) {
      double* out2 = A.lfac + (size_t)(2 * kp + 1) * kBlk * kBlk;
      for (int l = t; l < kBlk * kBlk; l += 64) out2[l] = sm.d2.D2[l >> 5][l & 31];
      for (int l = t; l < kBlk * kBlk; l += 64) {
        const int r = l >> 5, k = l & 31;
        if (r < nb2 && k < nb1) A.sys[(size_t)(c0 + kBlk + r) * ld + c0 + k] = sm.L21[r][k];
      }
    }
  }
  if (tid < kPanelRows / 64) {
    const int g = blockIdx.x * (kPanelRows / 64) + tid;  // 64-row group below the panel
    A.pflag[g] = sm.nz[tid];
    // persistent copy for the back substitution: rows of absolute group kp + 1 + g have non-zeros in panel kp
    if (w == kPanW && kp + 1 + g < A.pstride) A.pmask[(size_t)kp * A.pstride + kp + 1 + g] = sm.nz[tid];
  }
}

// C[r][c] -= sum_k X[r][k] X[c][k] on 64x64 tiles (tr, tc), tc <= tr, of the trailing matrix that starts at
// t0 = c0 + 32; 256 threads, 4x4 outputs each.  A CTA walks its live tiles with a two-deep pipeline: the panel rows of
// the NEXT tile and the old C values of the current one are in flight (registers) while the current tile is multiplied
// out of shared memory, so a tile costs its arithmetic plus one shared-memory round trip instead of three dependent L2
// round trips (load panels -> compute -> read-modify-write C), which made a tile 6 us for 2 us of DFMAs (r1).
constexpr int kTileQ = kUpdTile * kPanW / 2 / kFusedThreads;  // double2 loads per thread for one 64 x 64 panel tile
struct TileRegs {
  double2 r[kTileQ], c[kTileQ];
};

__device__ __forceinline__ void tile_decode(int t, int& tr, int& tc) {
  tr = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
  while ((tr + 1) * (tr + 2) / 2 <= t) tr++;
  while (tr * (tr + 1) / 2 > t) tr--;
  tc = t - tr * (tr + 1) / 2;
}

__device__ __forceinline__ void tile_load(const FusedArgs& A, int c0, int tr, int tc, TileRegs& g) {
  const int t0 = c0 + kPanW;
  const int r_base = t0 + tr * kUpdTile, c_base = t0 + tc * kUpdTile;
#pragma unroll
  for (int q = 0; q < kTileQ; q++) {
    const int l = threadIdx.x + q * kFusedThreads;
    const int r = l / (kPanW / 2), k2 = l % (kPanW / 2);
    const int gr = r_base + r, gc = c_base + r;
    g.r[q] = make_double2(0.0, 0.0);
    g.c[q] = make_double2(0.0, 0.0);
    if (gr <= A.n) g.r[q] = __ldcg(reinterpret_cast<const double2*>(A.sys + (size_t)gr * A.ld + c0) + k2);
    if (gc < A.n) g.c[q] = __ldcg(reinterpret_cast<const double2*>(A.sys + (size_t)gc * A.ld + c0) + k2);
  }
}

__device__ __forceinline__ void tile_store(UpdateSmem& sm, const TileRegs& g) {
#pragma unroll
  for (int q = 0; q < kTileQ; q++) {
    const int l = threadIdx.x + q * kFusedThreads;
    const int r = l / (kPanW / 2), k2 = l % (kPanW / 2);
    sm.R[r][2 * k2] = g.r[q].x;
    sm.R[r][2 * k2 + 1] = g.r[q].y;
    sm.C[r][2 * k2] = g.c[q].x;
    sm.C[r][2 * k2 + 1] = g.c[q].y;
  }
}

__device__ __forceinline__ void update_phase(const FusedArgs& A, int kp, int nfac, UpdateSmem& sm) {
  const int n = A.n, ld = A.ld;
  const int c0 = kp * kPanW, t0 = c0 + kPanW;
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
    for (int k = 0; k < kPanW; k++) {
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
  double L[2][kBlk][kBlk + 2];  // L_kk, double-buffered: the next block's factor arrives (cp.async) during this block
  double x[kBlk];
  double norm[kFusedThreads / 32];
  int live[64];                 // column panels with non-zeros in the current block row
  int fb[kSkyMaxBlk];           // envelope solver: first non-zero column block of every row block
  int bad;
};

__device__ __forceinline__ void back_stage(const FusedArgs& A, BackSmem& sm, int kb) {
  const double* src = A.lfac + (size_t)kb * kBlk * kBlk;
  for (int l = threadIdx.x; l < kBlk * kBlk / 2; l += kFusedThreads)
    cp_async16_zfill(&sm.L[kb & 1][l >> 4][2 * (l & 15)], src + 2 * l, 16);
  asm volatile("cp.async.commit_group;" ::: "memory");
}

// History (n = 1785, r1): 405 us with L_kk fetched on the critical path and the y update running 8 loads deep over
// every column; now the next factor is prefetched, the update has all 32 loads of a column in flight and skips column
// panels that the factorisation flagged as zero for this block row (pmask).
// SKY: the envelope solver ran; s_y already holds L^-1 b, and block row kb of L is non-zero in column blocks
// [fb[kb], kb) only.
template <bool SKY>
__device__ __forceinline__ void backsolve_phase(const FusedArgs& A, BackSmem& sm, double* s_y) {
  const int n = A.n, ld = A.ld, nblk = A.nblk;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) sm.bad = __ldcg(A.state + kFailNow);
  if (SKY) {
    for (int i = tid; i < nblk; i += kFusedThreads) sm.fb[i] = __ldcg(A.sky + SkyPlan::fb(A.n_free) + i);
  } else {
    for (int c = tid; c < nblk * kBlk; c += kFusedThreads) s_y[c] = (c < n) ? __ldcg(A.sys + (size_t)n * ld + c) : 0.0;
  }
  back_stage(A, sm, nblk - 1);
  for (int kb = nblk - 1; kb >= 0; kb--) {
    const int c0 = kb * kBlk;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();  // L_kk landed for everyone; the previous block's y update is complete
    if (kb > 0) back_stage(A, sm, kb - 1);
    const double (*L)[kBlk + 2] = sm.L[kb & 1];
    const int rows = min(kBlk, n - c0);
    // envelope solver: a block row has few columns (one per thread or less, except the long rows), and they do not
    // depend on x -- fetch this thread's first column now, so that its L2 round trip overlaps the triangular solve
    const int cfirst = (SKY ? sm.fb[kb] * kBlk : 0) + tid;
    double vp[kBlk];
    if (SKY && cfirst < c0) {
#pragma unroll
      for (int r = 0; r < kBlk; r++) vp[r] = r < rows ? __ldcg(A.sys + (size_t)(c0 + r) * ld + cfirst) : 0.0;
    }
    if (warp == 0) {
      double yv = s_y[c0 + lane];
      const double dinv = 1.0 / L[lane][lane];
#pragma unroll
      for (int j = kBlk - 1; j >= 0; j--) {
        const double xj = shfl_d(yv * dinv, j);  // lane j holds the finished y_j
        if (lane == j) yv = xj;
        if (lane < j) yv -= L[j][lane] * xj;
      }
      sm.x[lane] = yv;
      s_y[c0 + lane] = yv;
    } else if (warp == 1 && !SKY) {
      // live column panels of this block row: flagged panels left of the diagonal group, plus the diagonal group's own
      // first half (L21) when this is the second 32-row block of its panel
      const int gb = c0 / kPanW;
      for (int q = lane; q < 64; q += 32) sm.live[q] = (q < gb) ? __ldcg(A.pmask + (size_t)q * A.pstride + gb) : 0;
    }
    __syncthreads();
    // y[0:c0] -= L[c0:c0+rows, 0:c0]^T x : one column per thread, coalesced along the row
    if (SKY && cfirst < c0) {
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int r = 0; r < kBlk; r += 2) {
        a0 = fma(vp[r], sm.x[r], a0);
        a1 = fma(vp[r + 1], sm.x[r + 1], a1);
      }
      s_y[cfirst] -= a0 + a1;
    }
    for (int c = cfirst + (SKY ? kFusedThreads : 0); c < c0; c += kFusedThreads) {
      const int q = c / kPanW;
      if (!SKY && q * kPanW + kPanW <= c0 && !sm.live[q & 63]) continue;  // zero panel (the diagonal group is never skipped)
      double v[kBlk];
#pragma unroll
      for (int r = 0; r < kBlk; r++) v[r] = r < rows ? __ldcg(A.sys + (size_t)(c0 + r) * ld + c) : 0.0;
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int r = 0; r < kBlk; r += 2) {
        a0 = fma(v[r], sm.x[r], a0);
        a1 = fma(v[r + 1], sm.x[r + 1], a1);
      }
      s_y[c] -= a0 + a1;
    }
  }
  __syncthreads();
  finish_solution(A, s_y, &sm.bad, sm.norm);
}

// ============================================================================================== envelope (skyline) solver
// Keyframe graphs of a SLAM run are chains with a few long-range edges: the assembled H is block-banded plus a few long
// rows, and a Cholesky factor keeps the row envelope of the matrix it factors.  When at most kSkyMaxLanes 32-row blocks
// are ever active at once (SkyPlan, gn_solve.cu) ONE CTA factors the whole system out of shared memory: the active tiles
// (row block x column block, both active) stay resident from the step a row block joins until its column has been
// eliminated, so each entry of the envelope is read from L2 once and written once; there is no grid barrier, no
// trailing-matrix round trip through L2 and no work on structural zeros.  The dense multi-CTA solver above spends
// 27 us per 64-column panel on the diagonal block's serial chain plus two grid barriers whatever the sparsity
// (1.28 ms at n = 1785); this one spends the chain alone.  The right-hand side travels as a vector in shared memory
// (forward substitution block by block), not as row n of the matrix, so it does not widen the envelope.
// Per step kb (column block c0 = 32 kb): join (cp.async the tiles of row blocks that become active) | factor the
// diagonal tile unless the look-ahead already did | X = X L^-T on the active rows below, y_kb = L^-1 y_kb | publish
// L_kk, update the next diagonal tile | warp 0 factors the NEXT diagonal tile while warps 1-7 store the finished column
// to sys, update y and the remaining active tiles.
// ---- TMA bulk copies (cp.async.bulk, 1-D) + mbarrier: a 32 x 32 fp64 tile moves as 32 rows of 256 bytes, one
//      instruction per lane, instead of 16 cp.async / 32 STG per lane with their address arithmetic (the update warps
//      spent 2.8 us per step issuing the prefetch and 1.9 us storing the finished column of L that way; r1)
constexpr int kSkyTiles = kSkyMaxLanes * (kSkyMaxLanes + 1) / 2;
constexpr int kSkyPrefetched = 1 << 8;  // schedule word 1: the new members' tiles were fetched during the previous step
constexpr int kSkyStride = kBlk + 2;  // 272-byte rows: conflict-free LDS.128 for "one row per lane" and "8 rows per quarter warp"
struct SkySmem {
  double T[kSkyTiles][kBlk][kSkyStride];
  DiagAux d;
  unsigned long long mbar;  // completion of the tile loads of one join (one arrival carrying the byte count)
  int lane[kSkyMaxBlk];
  int adj[kSkyMaxBlk][4];  // bit lo of row hi (lo <= hi): the tile of row blocks {hi, lo} has a structural non-zero to load
  int sched[kSkyMaxBlk][kSkySched];  // the whole plan: no L2 round trip on the per-step critical path
  int order[kSkyMaxBlk];   // segmented solver: this stage's pivots in elimination order
};


__device__ __forceinline__ double (*sky_tile(SkySmem& sm, int la, int lb))[kSkyStride] {
  const int p = max(la, lb), q = min(la, lb);
  return sm.T[p * (p + 1) / 2 + q];
}

// C[16 half .. +16)[0..32) -= Xi Xj^T (K = 32) by one warp: 4 x 8 threads, 4 x 4 outputs each (rows 16 half + 4 ty + i,
// columns tx + 8 j: the eight lanes of a quarter warp read eight consecutive rows of Xj, one row of Xi)
__device__ __noinline__ void sky_update_item(const double (*Xi)[kSkyStride], const double (*Xj)[kSkyStride],
                                                double (*Cc)[kSkyStride], int half, int lane) {
  const int ty = lane >> 3, tx = lane & 7;
  const int r0 = 16 * half + 4 * ty;
  double acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j] = 0.0;
#pragma unroll 4
  for (int k = 0; k < kBlk; k += 2) {
    double2 a[4], b[4];
#pragma unroll
    for (int i = 0; i < 4; i++) a[i] = *reinterpret_cast<const double2*>(&Xi[r0 + i][k]);
#pragma unroll
    for (int j = 0; j < 4; j++) b[j] = *reinterpret_cast<const double2*>(&Xj[tx + 8 * j][k]);
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) acc[i][j] = fma(a[i].y, b[j].y, fma(a[i].x, b[j].x, acc[i][j]));
  }
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) Cc[r0 + i][tx + 8 * j] -= acc[i][j];
}

// One warp: right-hand side block of row block rbj, y -= X_j y_kb (one row per lane), and the finished column block of L
// (rows of block rbj, columns c0 .. c0 + nb) to sys for the back substitution
__device__ __forceinline__ void sky_rhs_and_store(const FusedArgs& A, SkySmem& sm, double* s_y, int rbj, int al0, int c0,
                                                  int nb, int lane) {
  const double (*Xj)[kSkyStride] = sky_tile(sm, sm.lane[rbj], al0);
  double s0 = 0.0, s1 = 0.0;
#pragma unroll
  for (int k = 0; k < kBlk; k += 2) {
    const double2 a = *reinterpret_cast<const double2*>(&Xj[lane][k]);
    s0 = fma(a.x, s_y[c0 + k], s0);
    s1 = fma(a.y, s_y[c0 + k + 1], s1);
  }
  s_y[kBlk * rbj + lane] -= s0 + s1;
  const int row = kBlk * rbj + lane;
  if (nb == kBlk) {  // always, when there are rows below: one 256-byte bulk store per row
    if (row < A.n) bulk_s2g(A.sys + (size_t)row * A.ld + c0, &Xj[lane][0], kBlk * sizeof(double));
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  } else {
    for (int r = 0; r < kBlk; r++) {
      const int rr = kBlk * rbj + r;
      if (rr < A.n && lane < nb) A.sys[(size_t)rr * A.ld + c0 + lane] = Xj[r][lane];
    }
  }
}

// Threads t0, t0 + 1, ... < nthr (whole warps): the tiles that join at step kb.  Full tiles: one 256-byte bulk copy per tile
// row, the (tile, row) pairs dealt out over the threads, completing on sm.mbar (thread 0 of the group posts the byte
// count and is the one arrival); rows beyond the matrix are zeroed with plain stores; the one tile with a partial
// column block (the last diagonal tile) uses zero-filling cp.async.  Ends with a cp.async commit.
__device__ __forceinline__ void sky_join(const FusedArgs& A, SkySmem& sm, int kb, int t0, int nthr) {
  const int n = A.n, ld = A.ld;
  const int* sched = sm.sched[kb];
  const int m = sched[0], mask = sched[1];
  unsigned bytes = 0;
  int q = 0;  // running (tile, row) index over the full tiles
  // pairs (new member p, any member o); a pair of two new members is taken once (by the larger index)
  for (int p = 0; p < m; p++) {
    if (!((mask >> p) & 1)) continue;
    for (int o = 0; o < m; o++) {
      if (o < p && ((mask >> o) & 1)) continue;
      const int i = max(p, o), j = min(p, o);
      const int rbi = sched[2 + i], rbj = sched[2 + j];
      double (*Tt)[kSkyStride] = sky_tile(sm, sm.lane[rbi], sm.lane[rbj]);
      const int hi = max(rbi, rbj), lo = min(rbi, rbj);
      if (!((sm.adj[hi][lo >> 5] >> (lo & 31)) & 1)) {
        // a tile inside the envelope that is structurally zero in the assembled matrix (it only receives fill): no copy
        for (int piece = t0; piece < 512; piece += nthr)
          *reinterpret_cast<double2*>(&Tt[piece >> 4][(piece & 15) * 2]) = make_double2(0.0, 0.0);
      } else if (rbi < rbj) {
        // segmented solver: the row block precedes the column block in the matrix (a separator in front of its interior);
        // the assembled tile is the transpose of (rbj, rbi) in the lower triangle.  Plain loads, a few tiles per interior.
        for (int piece = t0; piece < kBlk * kBlk; piece += nthr) {
          const int r = piece >> 5, c = piece & 31;  // source element (row r of block rbj, column c of block rbi)
          const int row = kBlk * rbj + r, col = kBlk * rbi + c;
          Tt[c][r] = (row < n && col < n) ? __ldcg(A.sys + (size_t)row * ld + col) : 0.0;
        }
      } else if (kBlk * rbj + kBlk <= n) {
        const int rows = min(kBlk, n - kBlk * rbi);
        bytes += rows * (unsigned)(kBlk * sizeof(double));
        // this thread's row of this tile, if any: pair index q + r with (q + r) % nthr == t0
        int r = (t0 - q % nthr + nthr) % nthr;
        for (; r < kBlk; r += nthr) {
          if (r < rows) {
            bulk_g2s(&Tt[r][0], A.sys + (size_t)(kBlk * rbi + r) * ld + kBlk * rbj, kBlk * sizeof(double), &sm.mbar);
          } else {
  #pragma unroll
            for (int c = 0; c < kBlk; c += 2) *reinterpret_cast<double2*>(&Tt[r][c]) = make_double2(0.0, 0.0);
          }
        }
        q += kBlk;
      } else {
        for (int piece = t0; piece < 512; piece += nthr) {  // 512 16-byte pieces
          const int r = piece >> 4, c2 = (piece & 15) * 2;
          const int row = kBlk * rbi + r, col = kBlk * rbj + c2;
          int cnt = (row < n) ? n - col : 0;
          cnt = cnt < 0 ? 0 : (cnt > 2 ? 2 : cnt);
          const double* src = cnt > 0 ? A.sys + (size_t)row * ld + col : A.sys;
          cp_async16_zfill(&Tt[r][c2], src, 8 * cnt);
        }
      }
    }
  }
  if (t0 == 0) mbar_arrive_expect_tx(&sm.mbar, bytes);
  asm volatile("cp.async.commit_group;" ::: "memory");
}

// all threads: the tiles of the pending join have landed (then a CTA barrier makes them visible to everybody)
__device__ __forceinline__ void sky_join_wait(const FusedArgs& A, SkySmem& sm, unsigned& parity) {
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  int spins = 0;
  while (!mbar_try_wait(&sm.mbar, parity)) {
    if (++spins > (1 << 24)) {  // never in a correct run: report a failed factorisation instead of hanging the device
      A.state[kFailNow] = 1;
      break;
    }
  }
  parity ^= 1u;
  __syncthreads();
}

// ND: stage `stage` of the segmented schedule (nd_plan.cuh) -- interior stage < ndP: the pivots of that interior, the
// separator tiles start from zero and are exported at the end; stage == ndP: the separators, s_y prepared by nd_reduce.
template <bool ND>
__device__ __forceinline__ void sky_factor(const FusedArgs& A, SkySmem& sm, double* s_y, int stage, int ndP) {
  const int n = A.n, ld = A.ld, nblk = A.nblk;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int cnt = nblk;
  if (!ND) {
    for (int i = tid; i < nblk; i += kFusedThreads) sm.lane[i] = __ldcg(A.sky + SkyPlan::lane(A.n_free) + i);
    for (int i = tid; i < nblk * kSkySched; i += kFusedThreads)
      sm.sched[i / kSkySched][i % kSkySched] = __ldcg(A.sky + SkyPlan::sched(A.n_free) + i);
    for (int i = tid; i < nblk * 4; i += kFusedThreads) sm.adj[i >> 2][i & 3] = __ldcg(A.sky + SkyPlan::adj(A.n_free) + i);
    for (int c = tid; c < nblk * kBlk; c += kFusedThreads) s_y[c] = (c < n) ? __ldcg(A.sys + (size_t)n * ld + c) : 0.0;
  } else {
    const int beg = __ldcg(A.nd + NdPlan::stage_begin + stage);
    cnt = __ldcg(A.nd + NdPlan::stage_begin + stage + 1) - beg;
    for (int i = tid; i < cnt; i += kFusedThreads) sm.order[i] = __ldcg(A.nd + NdPlan::order + beg + i);
    for (int i = tid; i < nblk; i += kFusedThreads) sm.lane[i] = __ldcg(A.nd + NdPlan::lane + stage * kSkyMaxBlk + i);
    for (int i = tid; i < nblk * kSkySched; i += kFusedThreads)
      sm.sched[i / kSkySched][i % kSkySched] = __ldcg(A.nd + NdPlan::sched + i);
    // tiles to load: an interior sees the assembled matrix without the separator-separator tiles (its Schur complement
    // starts from zero), the separator stage sees them with the fill of the interiors
    for (int i = tid; i < nblk * 4; i += kFusedThreads) {
      const int rb = i >> 2, w = i & 3;
      const unsigned in_s = (unsigned)__ldcg(A.nd + NdPlan::inS + w);
      const bool row_s = ((unsigned)__ldcg(A.nd + NdPlan::inS + (rb >> 5)) >> (rb & 31)) & 1u;
      unsigned v;
      if (stage == ndP) {
        v = (unsigned)__ldcg(A.nd + NdPlan::adjS + i);
      } else {
        v = (unsigned)__ldcg(A.sky + SkyPlan::adj(A.n_free) + i);
        if (row_s) v &= ~in_s;
      }
      sm.adj[rb][w] = (int)v;
    }
    if (stage < ndP)
      for (int c = tid; c < nblk * kBlk; c += kFusedThreads)
        s_y[c] = (c < n && __ldcg(A.nd + NdPlan::stage_of + c / kBlk) == stage) ? __ldcg(A.sys + (size_t)n * ld + c) : 0.0;
  }
  if (tid == 0) mbar_init(&sm.mbar, 1);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  unsigned jparity = 0;  // phase of sm.mbar the next join completes (uniform over the CTA)
  bool looked = false;   // the diagonal tile of this step was factored during the previous step (uniform over the CTA)
  for (int step = 0; step < cnt; step++) {
    const int kb = ND ? sm.order[step] : step;                                    // pivot block
    const int kn = step + 1 < cnt ? (ND ? sm.order[step + 1] : step + 1) : -1;  // next pivot of this stage
    const int c0 = kb * kBlk;
    const int nb = min(kBlk, n - c0);
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // my bulk stores of the last column have read their tiles
    __syncthreads();  // everybody is done with the previous step (the tiles of its column); first step: plan loaded
#define SKY_MARK(t) if (kb == 20 || kb == 21) tl_mark(A, (t))
    SKY_MARK(7000);
    const int* sched = sm.sched[kb];
    const int m = sched[0], mask = sched[1];
    const int al0 = sm.lane[kb];
    // ---- join: tiles (act[i], act[j]), j <= i, with a new member on either side, from the assembled system; when the
    //      plan says that the lanes of the NEXT step's new members are idle during this step, their tiles are fetched now
    if (mask & 0xff) {
      if (!(mask & kSkyPrefetched)) sky_join(A, sm, kb, tid, kFusedThreads);
      sky_join_wait(A, sm, jparity);
    }
    SKY_MARK(7001);
    double (*Tkk)[kSkyStride] = sky_tile(sm, al0, al0);
    if (!looked) {
      if (warp == 0) {
        const bool bad = factor_diag32(&Tkk[lane][0], sm.d, nullptr, nb, lane);
        if (bad && lane == 0) A.state[kFailNow] = 1;
      }
      __syncthreads();
    }
    // ---- rows below: X = X L_kk^-T, one row per thread (warp w: row block act[1 + w]); lane 0 of the first free warp
    //      forward-substitutes the right-hand side block
    if (warp < m - 1) {
      trsm_row32(&sky_tile(sm, sm.lane[sched[3 + warp]], al0)[lane][0], sm.d, nb);
    } else if (tid == (m - 1) * 32) {
      trsm_row32(s_y + c0, sm.d, nb);
    }
    // the bulk stores of this column (async proxy) read what was just written: writer-side proxy fence, then a CTA barrier,
    // then the stores.  With the look-ahead the fence waits until after the diagonal update (0.4 us on the critical path
    // right here, nothing there: the writes have long drained)
    const bool ahead = m >= 2 && sched[3] == kn;
    if (!ahead) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    SKY_MARK(7003);
    // ---- update of the next diagonal tile first (look-ahead): thread = (4 rows, 1 column)
    const int al1 = sm.lane[kn >= 0 ? kn : kb];
    if (ahead) {
      const double (*X1)[kSkyStride] = sky_tile(sm, al1, al0);
      double (*C11)[kSkyStride] = sky_tile(sm, al1, al1);
      const int r0 = 4 * warp;
      if (lane <= r0 + 3) {
        double s[4] = {0.0, 0.0, 0.0, 0.0}, u[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 4
        for (int k = 0; k < kBlk; k += 2) {
          const double2 b = *reinterpret_cast<const double2*>(&X1[lane][k]);
#pragma unroll
          for (int i = 0; i < 4; i++) {
            const double2 a = *reinterpret_cast<const double2*>(&X1[r0 + i][k]);
            s[i] = fma(a.x, b.x, s[i]);
            u[i] = fma(a.y, b.y, u[i]);
          }
        }
#pragma unroll
        for (int i = 0; i < 4; i++) C11[r0 + i][lane] -= s[i] + u[i];
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
    }
    SKY_MARK(7004);
    if (ahead && warp == 0) {
      const int c1 = kn * kBlk;
      const bool bad = factor_diag32(&sky_tile(sm, al1, al1)[lane][0], sm.d, nullptr, min(kBlk, n - c1), lane);
      if (bad && lane == 0) A.state[kFailNow] = 1;
      SKY_MARK(7005);
    } else {
      // workers: L_kk for the back substitution (zero upper part, identity padding) from the factored diagonal tile, the
      // right-hand side and the finished column of L, the prefetch of the next step's tiles, then the trailing update
      // with the look-ahead warp 4 is numbered last: it shares its scheduler with the factoring warp 0 and so gets the
      // fewest update items
      const int worker = ahead ? (warp < 4 ? warp - 1 : (warp == 4 ? kFusedThreads / 32 - 2 : warp - 2)) : warp;
      const int nworkers = ahead ? kFusedThreads / 32 - 1 : kFusedThreads / 32;
      const int pairs = m * (m - 1) / 2, skip = ahead ? 1 : 0;
      const int nitems = 2 * (pairs - skip);
      // the first nitems % nworkers warps get one update item more than the others; the others ("light" warps) take
      // the remaining duties, so that no update warp outlasts the factoring warp (the 3-item warps also stored, solved
      // the right-hand side and issued loads before: 8.3 us against 6.0 us for the factorisation; r1)
      const int w0 = nitems % nworkers;
      const int nlight = nworkers - w0;
      if (worker >= w0) {
        const int lw = worker - w0;
        double* out = A.lfac + (size_t)kb * kBlk * kBlk;
        for (int l = lw * 32 + lane; l < kBlk * kBlk; l += nlight * 32) {
          const int r = l >> 5, c = l & 31;
          out[l] = (r < nb && c <= r) ? Tkk[r][c] : (r == c ? 1.0 : 0.0);
        }
        if (kb == 20 && tid == 224) tl_mark_any(A, 7101);
        for (int j = 1 + lw; j < m; j += nlight) sky_rhs_and_store(A, sm, s_y, sched[2 + j], al0, c0, nb, lane);
        if (kb == 20 && tid == 224) tl_mark_any(A, 7102);
        if (kn >= 0 && (sm.sched[kn][1] & kSkyPrefetched)) sky_join(A, sm, kn, lw * 32 + lane, nlight * 32);
        if (kb == 20 && tid == 224) tl_mark_any(A, 7103);
      }
      // trailing tiles (act[i], act[j]) -= X_i X_j^T, 1 <= j <= i < m, two 16-row halves each; (1, 1) is done above
      for (int item = worker; item < nitems; item += nworkers) {
        int ti, tj;
        tile_decode((item >> 1) + skip, ti, tj);
        const int ali = sm.lane[sched[3 + ti]], alj = sm.lane[sched[3 + tj]];
        sky_update_item(sky_tile(sm, ali, al0), sky_tile(sm, alj, al0), sky_tile(sm, ali, alj), item & 1, lane);
        if (kb == 20 && tid == 224) tl_mark_any(A, 7110 + item);
      }
    }
    looked = ahead;
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // every column of L has reached global memory
  __syncthreads();
  if (ND && stage < ndP) {
    // the interior is eliminated; what its steps accumulated on the border blocks is its Schur complement: tiles
    // (border i, border j), j <= i, and the border parts of the right-hand side, to this interior's scratch slot
    const int* ex = A.nd + NdPlan::exp + stage * (1 + kNdMaxBorder);
    const int nbr = __ldcg(ex);
    double* dst = A.ndbuf + NdBuf::schur + (size_t)stage * kNdBorderTiles * kBlk * kBlk;
    for (int i = 0; i < nbr; i++)
      for (int j = 0; j <= i; j++) {
        const double (*Tt)[kSkyStride] = sky_tile(sm, sm.lane[__ldcg(ex + 1 + i)], sm.lane[__ldcg(ex + 1 + j)]);
        double* d = dst + (size_t)(i * (i + 1) / 2 + j) * kBlk * kBlk;
        for (int e = tid; e < kBlk * kBlk; e += kFusedThreads) d[e] = Tt[e >> 5][e & 31];
      }
    if (tid < nbr * kBlk)
      A.ndbuf[NdBuf::srhs + ((size_t)stage * kNdMaxBorder + (tid >> 5)) * kBlk + lane] =
          s_y[kBlk * __ldcg(ex + 1 + (tid >> 5)) + lane];
  }
}

// ============================================================================================== segmented envelope solver
// Stage layout of the launch (nd_plan.cuh): CTA 0 owns the separator stage, CTA 1 + s interior s.  Three barriers over
// these 1 + P CTAs: interiors exported | separators solved | interiors solved.

// CTA 0: right-hand side and tiles of the separator system = assembled values + the interiors' Schur complements, summed
// in interior order by a fixed thread per element (deterministic: the replicas of a multi-GPU run stay bit-identical).
__device__ __forceinline__ void nd_reduce(const FusedArgs& A, double* s_y, int ndP, int* s_list /* [32 + kNdMaxSeg * 8] */) {
  const int n = A.n, ld = A.ld, nblk = A.nblk;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int* sblk = s_list;        // separator blocks, ascending; [31] = their number
  int* ex = s_list + 32;     // [ndP][1 + kNdMaxBorder] border lists
  unsigned in_s[4];
#pragma unroll
  for (int w = 0; w < 4; w++) in_s[w] = (unsigned)__ldcg(A.nd + NdPlan::inS + w);
  if (tid == 0) {
    int k = 0;
    for (int w = 0; w < 4; w++) {
      unsigned m = in_s[w];
      while (m && k < 31) {
        sblk[k++] = 32 * w + __ffs((int)m) - 1;
        m &= m - 1;
      }
    }
    sblk[31] = k;
  }
  for (int i = tid; i < ndP * (1 + kNdMaxBorder); i += kFusedThreads) ex[i] = __ldcg(A.nd + NdPlan::exp + i);
  __syncthreads();
  // right-hand side: a fixed thread per element, contributions added in interior order
  for (int c = tid; c < nblk * kBlk; c += kFusedThreads) {
    const int rb = c / kBlk;
    double v = 0.0;
    if (((in_s[rb >> 5] >> (rb & 31)) & 1u) && c < n) {
      double part[kNdMaxSeg];
#pragma unroll
      for (int s = 0; s < kNdMaxSeg; s++) {
        part[s] = 0.0;
        if (s < ndP) {
          const int nbr = ex[s * (1 + kNdMaxBorder)];
          for (int i = 0; i < nbr; i++)
            if (ex[s * (1 + kNdMaxBorder) + 1 + i] == rb)
              part[s] = __ldcg(A.ndbuf + NdBuf::srhs + ((size_t)s * kNdMaxBorder + i) * kBlk + (c & 31));
        }
      }
      v = __ldcg(A.sys + (size_t)n * ld + c);
#pragma unroll
      for (int s = 0; s < kNdMaxSeg; s++) v += part[s];  // adding 0.0 for an interior that does not touch the block is exact
    }
    s_y[c] = v;
  }
  // tiles: one warp per pair of separator blocks (a >= b) that the separator stage will load; eight rows in flight
  const int ns = sblk[31];
  for (int pr = warp; pr < ns * (ns + 1) / 2; pr += kFusedThreads / 32) {
    int ia, ib;
    tile_decode(pr, ia, ib);
    const int a = sblk[ia], b = sblk[ib];
    if (!(((unsigned)__ldcg(A.nd + NdPlan::adjS + a * 4 + (b >> 5)) >> (b & 31)) & 1u)) continue;
    const bool assembled = ((unsigned)__ldcg(A.sky + SkyPlan::adj(A.n_free) + a * 4 + (b >> 5)) >> (b & 31)) & 1u;
    // sources: interiors whose border list holds both blocks
    int src[kNdMaxSeg], nsrc = 0;
    for (int s = 0; s < ndP; s++) {
      const int nbr = ex[s * (1 + kNdMaxBorder)];
      int pa = -1, pb = -1;
      for (int i = 0; i < nbr; i++) {
        const int blk = ex[s * (1 + kNdMaxBorder) + 1 + i];
        if (blk == a) pa = i;
        if (blk == b) pb = i;
      }
      if (pa >= 0 && pb >= 0) src[nsrc++] = (s * kNdBorderTiles + pa * (pa + 1) / 2 + pb);
    }
    // no source and assembled: sys already holds the tile.  No source and NOT assembled: the tile only receives fill
    // inside the separator stage, but that stage loads every tile of its plan from sys -- which, outside the envelope that
    // the assembly zeroes, still holds the factor of the previous iteration: write the zeros.
    if (nsrc == 0 && assembled) continue;
    const int col = kBlk * b + lane;
    for (int r0 = 0; r0 < kBlk; r0 += 8) {
      double v[8];
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const int row = kBlk * a + r0 + i;
        v[i] = (assembled && row < n && col < n) ? __ldcg(A.sys + (size_t)row * ld + col) : 0.0;
      }
      for (int q = 0; q < nsrc; q++) {
        double t[8];
#pragma unroll
        for (int i = 0; i < 8; i++) t[i] = __ldcg(A.ndbuf + NdBuf::schur + (size_t)src[q] * kBlk * kBlk + (r0 + i) * kBlk + lane);
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] += t[i];
      }
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const int row = kBlk * a + r0 + i;
        if (row < n && col < n && !(a == b && lane > r0 + i)) A.sys[(size_t)row * ld + col] = v[i];
      }
    }
  }
  asm volatile("fence.proxy.async.global;" ::: "memory");  // the separator stage's joins read these through the copy engine
  __syncthreads();
}

// Back substitution over the block rows rows[0 .. cnt) in that order.  SOLVE: x_kb = L_kk^-T y_kb first (the rows of a
// stage, last pivot first); otherwise s_y already holds x_kb (an interior's border blocks).  Then y[c] -= L(kb, c)^T x_kb
// for the column blocks of L's block row kb (NdPlan::lrow) that lie in `cols` -- wherever the tile sits in sys.
struct NdBackSmem {
  BackSmem b;
  unsigned lrow[kSkyMaxBlk][4];
  unsigned cols[4];
  int rows[kSkyMaxBlk];
  int brows[kNdMaxBorder];
};

__device__ __forceinline__ void nd_back_stage(const FusedArgs& A, BackSmem& sm, int kb, int buf) {
  const double* src = A.lfac + (size_t)kb * kBlk * kBlk;
  for (int l = threadIdx.x; l < kBlk * kBlk / 2; l += kFusedThreads)
    cp_async16_zfill(&sm.L[buf][l >> 4][2 * (l & 15)], src + 2 * l, 16);
  asm volatile("cp.async.commit_group;" ::: "memory");
}

template <bool SOLVE>
__device__ __forceinline__ void nd_back_rows(const FusedArgs& A, NdBackSmem& sm, double* s_y, const int* rowlist, int cnt) {
  const int n = A.n, ld = A.ld;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int kWarps = kFusedThreads / 32;
  if (cnt <= 0) return;
  if (SOLVE) nd_back_stage(A, sm.b, rowlist[0], 0);
  for (int i = 0; i < cnt; i++) {
    const int kb = rowlist[i];
    const int c0 = kb * kBlk;
    const int rows = min(kBlk, n - c0);
    // column blocks of this block row, dealt round-robin over the warps; the warp's FIRST one does not depend on x_kb:
    // it is fetched now, so that its L2 round trip overlaps the triangular solve below
    unsigned m[4];
#pragma unroll
    for (int w = 0; w < 4; w++) m[w] = sm.lrow[kb][w] & sm.cols[w];
    int first_cb = -1;
    {
      int slot = 0;
#pragma unroll
      for (int w = 0; w < 4; w++) {
        unsigned mm = m[w];
        while (mm && first_cb < 0) {
          const int cb = 32 * w + __ffs((int)mm) - 1;
          mm &= mm - 1;
          if (slot++ == warp) first_cb = cb;
        }
      }
    }
    double vp[kBlk];
    if (first_cb >= 0) {
      const int c = kBlk * first_cb + lane;
#pragma unroll
      for (int r = 0; r < kBlk; r++) vp[r] = (r < rows && c < n) ? __ldcg(A.sys + (size_t)(c0 + r) * ld + c) : 0.0;
    }
    if (SOLVE) {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();  // L_kk landed for everyone; the previous row's y update is complete
      if (i + 1 < cnt) nd_back_stage(A, sm.b, rowlist[i + 1], (i + 1) & 1);
      if (warp == 0) {
        const double (*L)[kBlk + 2] = sm.b.L[i & 1];
        double yv = s_y[c0 + lane];
        const double dinv = 1.0 / L[lane][lane];
#pragma unroll
        for (int j = kBlk - 1; j >= 0; j--) {
          const double xj = shfl_d(yv * dinv, j);  // lane j holds the finished y_j
          if (lane == j) yv = xj;
          if (lane < j) yv -= L[j][lane] * xj;
        }
        s_y[c0 + lane] = yv;
      }
    }
    __syncthreads();
    if (first_cb >= 0) {
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int r = 0; r < kBlk; r += 2) {
        a0 = fma(vp[r], s_y[c0 + r], a0);
        a1 = fma(vp[r + 1], s_y[c0 + r + 1], a1);
      }
      s_y[kBlk * first_cb + lane] -= a0 + a1;
    }
    // the warp's other column blocks (rows of L with more than kWarps of them: the long rows of a separator)
    int slot = 0;
    for (int w = 0; w < 4; w++) {
      unsigned mm = m[w];
      while (mm) {
        const int cb = 32 * w + __ffs((int)mm) - 1;
        mm &= mm - 1;
        const int sl = slot++;
        if (sl < kWarps || (sl & (kWarps - 1)) != warp) continue;
        const int c = kBlk * cb + lane;
        double v[kBlk];
#pragma unroll
        for (int r = 0; r < kBlk; r++) v[r] = (r < rows && c < n) ? __ldcg(A.sys + (size_t)(c0 + r) * ld + c) : 0.0;
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int r = 0; r < kBlk; r += 2) {
          a0 = fma(v[r], s_y[c0 + r], a0);
          a1 = fma(v[r + 1], s_y[c0 + r + 1], a1);
        }
        s_y[c] -= a0 + a1;
      }
    }
    if (!SOLVE) __syncthreads();
  }
  __syncthreads();
}

// stage tables of the back substitution: the rows of `stage` in reverse elimination order, the columns of that stage
__device__ __forceinline__ void nd_back_setup(const FusedArgs& A, NdBackSmem& sm, int stage, int& cnt) {
  const int tid = threadIdx.x;
  const int beg = __ldcg(A.nd + NdPlan::stage_begin + stage);
  cnt = __ldcg(A.nd + NdPlan::stage_begin + stage + 1) - beg;
  for (int i = tid; i < cnt; i += kFusedThreads) sm.rows[i] = __ldcg(A.nd + NdPlan::order + beg + cnt - 1 - i);
  for (int i = tid; i < A.nblk * 4; i += kFusedThreads) sm.lrow[i >> 2][i & 3] = (unsigned)__ldcg(A.nd + NdPlan::lrow + i);
  if (tid < 4) sm.cols[tid] = 0;
  __syncthreads();
  for (int rb = tid; rb < A.nblk; rb += kFusedThreads)
    if (__ldcg(A.nd + NdPlan::stage_of + rb) == stage) atomicOr(&sm.cols[rb >> 5], 1u << (rb & 31));
  __syncthreads();
}

union FusedSmem;
__device__ __forceinline__ void nd_solve(const FusedArgs& A, FusedSmem& sm, double* s_y, int ndP);

// ============================================================================================== the kernel
union FusedSmem {
  FinalizeScratch fin[kFusedThreads / 32];
  PanelSmem panel;
  UpdateSmem upd;
  BackSmem back;
  SkySmem sky;
  NdBackSmem ndback;
  int ndlist[32 + kNdMaxSeg * (1 + kNdMaxBorder)];
};

__device__ __forceinline__ void nd_solve(const FusedArgs& A, FusedSmem& sm, double* s_y, int ndP) {
  const int tid = threadIdx.x;
  const int nbar = ndP + 1;
  double* xg = A.ndbuf + NdBuf::x;
  if (blockIdx.x == 0) {
    // ---- separators
    grid_barrier(A.state, nbar);  // the interiors are eliminated, their Schur complements are in the scratch
    tl_mark(A, 8100);
    nd_reduce(A, s_y, ndP, sm.ndlist);
    tl_mark(A, 8200);
    sky_factor<true>(A, sm.sky, s_y, ndP, ndP);
    tl_mark(A, 8300);
    int cnt;
    nd_back_setup(A, sm.ndback, ndP, cnt);
    nd_back_rows<true>(A, sm.ndback, s_y, sm.ndback.rows, cnt);
    for (int i = tid; i < cnt * kBlk; i += kFusedThreads) {
      const int c = kBlk * sm.ndback.rows[i >> 5] + (i & 31);
      xg[c] = s_y[c];
    }
    tl_mark(A, 8400);
    grid_barrier(A.state, nbar);  // the separators' solution is out
    grid_barrier(A.state, nbar);  // the interiors' solutions are out
    tl_mark(A, 8500);
    for (int c = tid; c < A.nblk * kBlk; c += kFusedThreads) s_y[c] = __ldcg(xg + c);
    if (tid == 0) sm.ndback.b.bad = __ldcg(A.state + kFailNow);
    __syncthreads();
    finish_solution(A, s_y, &sm.ndback.b.bad, sm.ndback.b.norm);
    tl_mark(A, 6);
  } else {
    // ---- interior s
    const int s = blockIdx.x - 1;
    sky_factor<true>(A, sm.sky, s_y, s, ndP);
    grid_barrier(A.state, nbar);
    grid_barrier(A.state, nbar);
    int cnt;
    nd_back_setup(A, sm.ndback, s, cnt);
    // the border blocks' solution, then their share of this interior's right-hand side
    const int* ex = A.nd + NdPlan::exp + s * (1 + kNdMaxBorder);
    const int nbr = __ldcg(ex);
    if (tid < nbr * kBlk) {
      const int c = kBlk * __ldcg(ex + 1 + (tid >> 5)) + (tid & 31);
      s_y[c] = __ldcg(xg + c);
    }
    if (tid < nbr) sm.ndback.brows[tid] = __ldcg(ex + 1 + tid);
    __syncthreads();
    nd_back_rows<false>(A, sm.ndback, s_y, sm.ndback.brows, nbr);
    nd_back_rows<true>(A, sm.ndback, s_y, sm.ndback.rows, cnt);
    for (int i = tid; i < cnt * kBlk; i += kFusedThreads) {
      const int c = kBlk * sm.ndback.rows[i >> 5] + (i & 31);
      xg[c] = s_y[c];
    }
    grid_barrier(A.state, nbar);
  }
}

__global__ void __launch_bounds__(kFusedThreads, 1) gn_fused_kernel(const FusedArgs A) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  FusedSmem& sm = *reinterpret_cast<FusedSmem*>(s_raw);
  double* s_y = reinterpret_cast<double*>(s_raw + A.sy_off);
  // kDone is only written by the last phase of an earlier launch on the same stream: uniform over the grid
  if (__ldcg(A.state + kDone)) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // envelope solver: decided once per solve on the device (skyline_plan_kernel), the same for every CTA
  const int sky = (A.sky != nullptr) ? __ldcg(A.state + kSkyLanes) : 0;
  tl_mark(A, 1);

  if (A.phases & M3R_PH_FINALIZE) {
    const int nwarp = gridDim.x * (kFusedThreads / 32);
    for (int e = blockIdx.x * (kFusedThreads / 32) + warp; e < A.num_edges; e += nwarp)
      finalize_edge(e, lane, sm.fin[warp], A.Twc, A.partial, A.eoff, A.ie, A.je, A.blocks, A.num_edges, A.hess_grid);
  }
  if (A.phases & M3R_PH_ASSEMBLE) {
    if (blockIdx.x == 0 && threadIdx.x == 0) A.state[kFailNow] = 0;
    assemble_zero(A, sky);
    tl_mark(A, 2);
    grid_barrier(A.state, gridDim.x);  // also orders finalize -> assemble
    tl_mark(A, 3);
    assemble_items(A);
    tl_mark(A, 4);
  }
  if (A.peer.world > 1) {
    peer_allreduce_phase<kFusedThreads>(A, sky);  // its barriers also order assembly -> factorisation
    if (peer_timed_out(A)) return;
  } else if ((A.phases & M3R_PH_FACTOR) && (A.phases & M3R_PH_ASSEMBLE)) {
    grid_barrier(A.state, gridDim.x);
  }
  const int nd_p = (sky > 0 && A.nd != nullptr && (A.phases & (M3R_PH_FACTOR | M3R_PH_BACKSOLVE)) == (M3R_PH_FACTOR | M3R_PH_BACKSOLVE))
                       ? __ldcg(A.state + kNdSegs) : 0;
  if (nd_p > 0) {  // segmented envelope solver: factorisation, back substitution and retraction on 1 + P CTAs
    if ((int)blockIdx.x > nd_p) return;
    tl_mark(A, 5);
    FusedArgs B = A;  // the planner's winning candidate (gn_solve.cu)
    B.nd = A.nd + (size_t)__ldcg(A.state + kNdSlot) * NdPlan::total;
    nd_solve(B, sm, s_y, nd_p);
    return;
  }
  if ((A.phases & M3R_PH_FACTOR) && sky > 0) {
    if (blockIdx.x != 0) return;
    tl_mark(A, 5);
    sky_factor<false>(A, sm.sky, s_y, 0, 0);
    tl_mark(A, 8);
  } else if (A.phases & M3R_PH_FACTOR) {
    // only the CTAs that can get work in block column kb take part in its two barriers; the trailing matrix shrinks, so
    // the set is non-increasing and a CTA that drops out never comes back
    tl_mark(A, 5);
    const int npan = (A.n + kPanW - 1) / kPanW;
    for (int kp = 0; kp < npan; kp++) {
      const int below = A.n + 1 - min(A.n, (kp + 1) * kPanW);  // rows under the panel, rhs row included
      const int Tk = (below + kUpdTile - 1) / kUpdTile;
      int nfac = Tk * (Tk + 1) / 2;
      if (nfac > (int)gridDim.x) nfac = gridDim.x;
      if ((int)blockIdx.x >= nfac) return;
      panel_phase(A, kp, sm.panel);
      tl_mark(A, 1000 + kp);
      grid_barrier(A.state, nfac);
      tl_mark(A, 2000 + kp);
      if ((kp + 1) * kPanW < A.n) {
        update_phase(A, kp, nfac, sm.upd);
        tl_mark(A, 3000 + kp);
        grid_barrier(A.state, nfac);
        tl_mark(A, 4000 + kp);
      }
    }
  }
  if ((A.phases & M3R_PH_BACKSOLVE) && blockIdx.x == 0) {
    if (sky > 0)
      backsolve_phase<true>(A, sm.back, s_y);
    else
      backsolve_phase<false>(A, sm.back, s_y);
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
    assemble_zero(A, 0);
    tl_mark(A, 2);
    grid_barrier(A.state, gridDim.x);
    tl_mark(A, 3);
    assemble_items(A);
    tl_mark(A, 4);
  }
  if (A.peer.world > 1) {
    peer_allreduce_phase<kSmallThreads>(A, 0);  // its barriers also order assembly -> factorisation
    if (peer_timed_out(A)) return;
  } else if ((A.phases & M3R_PH_FACTOR) && (A.phases & M3R_PH_ASSEMBLE)) {
    grid_barrier(A.state, gridDim.x);
  }
  if (A.phases & M3R_PH_FACTOR) {
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
  if (want < kNdMaxSeg + 1) want = kNdMaxSeg + 1;  // the segmented envelope solver: one CTA per interior + one
  const int cap = sm_count();
  return want < cap ? want : cap;
}

int launch_fused(const GnWs& ws, float* Twc, float delta_thresh, int phases, cudaStream_t st, const PeerInfo* peer) {
  if (ws.L.n_free <= 0) return M3R_OK;
  FusedArgs a;
  a.peer = PeerInfo();
  a.peer.world = 1;
  if (peer != nullptr && peer->world > 1) a.peer = *peer;
  a.Twc = Twc;
  a.partial = ws.partial();
  a.eoff = ws.eoff();
  a.ie = ws.ie();
  a.je = ws.je();
  a.blocks = ws.blocks();
  a.inc_off = ws.inc_off();
  a.inc_list = ws.inc_list();
  a.inc_other = ws.inc_other();
  a.sys = a.peer.world > 1 ? a.peer.sys[a.peer.rank] : ws.sys();
  a.lfac = ws.lfac();
  a.pflag = ws.pflag();
  a.pmask = ws.pflag() + (ws.L.n / 64 + 4);
  a.pstride = ws.L.n / 64 + 4;
  a.sky = nullptr;
  a.nd = nullptr;
  a.ndbuf = nullptr;
  a.sy_off = 0;
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
    static DynSmemOptIn attr_small;
    {
      const int rc = ensure_dyn_smem(gn_small_kernel, smem, attr_small);
      if (rc != M3R_OK) return rc;
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
  // the envelope solver's tile store (183 KB) is only reserved when that solver can run; systems beyond its range keep
  // the dense solver's footprint (101 KB + the solution vector)
  a.sky = (g_solver_mode != 1 && ws.L.nblk <= kSkyMaxBlk) ? ws.sky() : nullptr;
  if (a.sky != nullptr && g_solver_mode != 2 && ws.L.off_nd != 0) {
    a.nd = ws.sky() + SkyPlan::total(ws.L.n_free);
    a.ndbuf = ws.ndbuf();
  }
  size_t phase_smem = sizeof(PanelSmem);
  if (sizeof(UpdateSmem) > phase_smem) phase_smem = sizeof(UpdateSmem);
  if (sizeof(BackSmem) > phase_smem) phase_smem = sizeof(BackSmem);
  if (sizeof(FinalizeScratch) * (kFusedThreads / 32) > phase_smem) phase_smem = sizeof(FinalizeScratch) * (kFusedThreads / 32);
  if (a.sky != nullptr) phase_smem = sizeof(FusedSmem);
  phase_smem = (phase_smem + 15) & ~(size_t)15;
  a.sy_off = (int)phase_smem;
  const size_t smem = phase_smem + sizeof(double) * (size_t)ws.L.nblk * kBlk;
  if (smem > 227 * 1024) {
    set_error("gauss_newton: system of dimension %d exceeds the shared-memory back substitution", ws.L.n);
    return M3R_ERR_UNSUPPORTED;
  }
  static DynSmemOptIn attr_smem;
  {
    const int rc = ensure_dyn_smem(gn_fused_kernel, smem, attr_smem);
    if (rc != M3R_OK) return rc;
  }
  M3R_CUDA(cudaLaunchCooperativeKernel((const void*)gn_fused_kernel, dim3((unsigned)fused_grid(ws)),
                                       dim3(kFusedThreads), args, smem, st));
  count_launch();
  return M3R_OK;
}

}  // namespace m3r

// diagnostics hook (tools/fused_timeline.py): device buffer of 2*cap u64; [0] = number of marks, then (tag, ns) pairs
extern "C" void m3r_debug_set_solver(int mode) { m3r::g_solver_mode = mode; }

extern "C" int m3r_debug_solver_lanes(const void* workspace) {
  int v = 0;  // the solver state is the first region of the workspace (gn_layout)
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  if (cudaMemcpy(&v, (const int*)workspace + m3r::kSkyLanes, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  return v;
}

extern "C" int m3r_debug_solver_segments(const void* workspace) {
  int v = 0;
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  if (cudaMemcpy(&v, (const int*)workspace + m3r::kNdSegs, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  return v;
}

extern "C" void m3r_debug_set_timeline(unsigned long long* dev_buf, int cap) {
  m3r::g_timeline = dev_buf;
  m3r::g_timeline_cap = cap;
}
