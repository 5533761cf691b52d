// Gauss-Newton Hessian assembly for sm_100a: per-edge Sim(3) residuals, Jacobians, IRLS weights and the reduction
// of J^T W J / J^T W r.  Replaces ray_align_kernel / calib_proj_kernel / point_align_kernel of the reference
// (gn_kernels.cu:813-1138, :1231-1543, :455-723) with a different decomposition:
//
//   * The reference accumulates the full 14x14 block per pixel with J_j = Adj(T_i)^-T J_local and J_i = -J_j
//     (gn_kernels.cu:999-1000).  Adj is constant per edge and linear, so we accumulate only
//     S = sum w J_local J_local^T (28 entries) and g = sum w e J_local (7 entries) in the local frame and apply
//     M = A S A^T, v = A g once per edge in fp64 (finalize kernel).  H_ii = H_jj = M, H_ij = H_ji = -M,
//     g_j = v, g_i = -v.
//   * Pose-independent work is hoisted into a packing pre-pass that runs once per solve: gates, the data-dependent
//     gathers of X_i / C_i and the normalisation of X_i end up in one self-contained 32-byte record per surviving
//     (pixel, edge); the per-iteration kernel streams those records with TMA bulk copies and gathers nothing.
//   * Persistent stream-K decomposition over the flat list of compacted records (the reference runs one 256-thread
//     block per edge), two records per thread in packed f32x2 arithmetic, fp32 warp-shuffle + shared-memory tree per
//     segment, deterministic fp64 second stage across segments (no atomics).
#include "f32x2.cuh"
#include "tma.cuh"
#include "gn_finalize.cuh"

namespace m3r {

// ------------------------------------------------------------------------------------------------ pre-pass
// Answers "does any point of any keyframe fail the confidence gate c > C_thresh?" (state[kCFail]); when none does -- the
// usual case, C_conf = 0 and MASt3R confidences are >= 1 -- pack_edges skips its two confidence gathers per pixel.
__global__ void __launch_bounds__(256) conf_scan_kernel(const float* __restrict__ Cs, long long total, float C_thresh,
                                                        int* __restrict__ cfail) {
  bool fail = false;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long long)gridDim.x * blockDim.x)
    fail |= !(__ldg(Cs + p) > C_thresh);  // NaN fails, as in the gate itself
  if (__any_sync(0xffffffffu, fail) && (threadIdx.x & 31) == 0) atomicOr(cfail, 1);
}

// Gate + compact + materialise.  One CTA per (edge, 8192-pixel tile); work items are handed out by a ticket so that the
// owners of the earlier tiles of an edge are guaranteed to be running (or finished) when a CTA looks back at them.
//   phase 1: every warp walks 1024 consecutive pixels (coalesced), evaluates the pose-independent gates and parks the
//            match index (or -1) in shared memory; survivors are counted with ballots.  The 8-byte index is only
//            fetched where valid_match and the Q gate pass, so whole sectors of it are never read outside the overlap
//            region of the two frames.
//   scan   : the CTA publishes its own count at once, then warp 0 sums the counts of the earlier tiles of the same edge
//            (decoupled look-back: nobody waits on a chain, only on each predecessor's phase 1).
//   phase 2: warps re-walk their pixels and write one SELF-CONTAINED 32-byte record per survivor, in pixel order, at the
//            scanned offset: everything the per-iteration kernel needs that does not depend on the poses -- the matched
//            point of frame i (already normalised for the ray residual), the point of frame j and sqrt(q).  The
//            data-dependent gathers Xs[ix][idx], Cs[ix][idx] (gn_kernels.cu:916-918, 951) are paid once per solve here
//            instead of once per iteration, and the Hessian kernel becomes a pure stream (TMA bulk copies, no gather).
// Where the 0.46 ms at configs[2] go (r2, kernel cut short after each phase): phase 1 0.18, scan + compaction 0.15,
// records 0.13 ms; DRAM traffic 1.06 GB = 2.3 TB/s; ncu: long scoreboard 12 and barrier 8 stall cycles per issue.  Tried and
// dropped: unrolling phase 1 1/4/8 x (0.538-0.555 ms, no change); a warp-independent version (a warp owns 1024 pixels
// from the gates to the records, one chained-scan slot per warp, no CTA barrier after the ticket): correct, bit-identical
// records, but SLOWER, 0.63 against 0.54 ms for the whole pre-pass -- eight times as many chained-scan slots to poll and
// 32 instead of 256 gathers in flight per round of the record phase cost more than the barriers did.
__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// record halves (GnRecord::a, GnRecord::b):
//   rays  : A = (r_i.x, r_i.y, r_i.z, |X_i|)  with r_i = X_i/|X_i|      B = (X_j.x, X_j.y, X_j.z, sqrt(q))
//   points: A = (X_i.x, X_i.y, X_i.z, 0)                                B = same
//   calib : A = (u_t, v_t, log z_i, 0)  target pixel of the match index (gn_kernels.cu:1361-1388)   B = same
__device__ __forceinline__ float4 make_rec_a(int kind, const float* __restrict__ Xi, int ind, int width) {
  const float x = __ldg(Xi + 3 * (size_t)ind), y = __ldg(Xi + 3 * (size_t)ind + 1), z = __ldg(Xi + 3 * (size_t)ind + 2);
  if (kind == M3R_GN_RAYS) {
    const float n = sqrtf(fmaf(x, x, fmaf(y, y, z * z)));
    const float inv = 1.0f / n;
    return make_float4(x * inv, y * inv, z * inv, n);
  }
  if (kind == M3R_GN_CALIB) {
    const int v_t = ind / width;
    return make_float4((float)(ind - v_t * width), (float)v_t, logf(z), 0.0f);
  }
  return make_float4(x, y, z, 0.0f);
}

constexpr size_t kPackSmemBytes = (size_t)kTilePix * (sizeof(int) + sizeof(unsigned short));
__global__ void __launch_bounds__(256) pack_edges_kernel(const float* __restrict__ Xs, const float* __restrict__ Cs,
                                                         const int64_t* __restrict__ idx,
                                                         const uint8_t* __restrict__ valid, const float* __restrict__ Q,
                                                         const int* __restrict__ ie, const int* __restrict__ je,
                                                         GnRecord* __restrict__ rec, int* __restrict__ ecount, int* __restrict__ tstate,
                                                         const int* __restrict__ cfail, int num_pts, int tiles,
                                                         float C_thresh, float Q_thresh, int kind, int width,
                                                         float z_eps) {
  // Every lane handles 4 consecutive pixels per pass (one LDG.32 of valid, one LDG.128 of Q, two predicated LDG.128 of
  // the 8-byte indices), 8 passes of 1024 pixels per tile: a quarter of the instructions of the one-pixel-per-lane
  // version, which was issue-bound at 52 % with 208 M warp instructions (r1).
  constexpr int kPasses = kTilePix / 1024;
  extern __shared__ __align__(16) int s_pack_dyn[];  // kPackSmemBytes
  int* s_idx = s_pack_dyn;                                                           // [kTilePix] match index or -1
  unsigned short* s_lk = reinterpret_cast<unsigned short*>(s_pack_dyn + kTilePix);  // [kTilePix] phase 2: pixel offset
  __shared__ int s_work, s_base, s_cnt[kPasses * 8];
  if (threadIdx.x == 0) s_work = atomicAdd(&tstate[0], 1);
  __syncthreads();
  const int work = s_work;
  const int e = work / tiles, tile = work - e * tiles;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ix = ie[e], jx = je[e];
  const bool edge_ok = ix >= 0 && jx >= 0;
  const bool cgate = __ldg(cfail) != 0;  // some point fails c > C_thresh: evaluate the confidence gates per pixel
  const bool zgate = kind == M3R_GN_CALIB;  // calib: z_i > z_eps is pose-independent (gn_kernels.cu:1368), gate it here
  const size_t eb = (size_t)e * num_pts;
  // vector path: every edge row of Q / idx starts 16-byte aligned and of valid 4-byte aligned
  const bool vec_ok = ((num_pts & 3) == 0) && ((reinterpret_cast<uintptr_t>(Q) & 15) == 0) &&
                      ((reinterpret_cast<uintptr_t>(idx) & 15) == 0) && ((reinterpret_cast<uintptr_t>(valid) & 3) == 0);
  const float* __restrict__ Ci = Cs + (size_t)(edge_ok ? ix : 0) * num_pts;
  const float* __restrict__ Cj = Cs + (size_t)(edge_ok ? jx : 0) * num_pts;
  const float* __restrict__ Xi = Xs + 3 * (size_t)(edge_ok ? ix : 0) * num_pts;
  const float* __restrict__ Xj = Xs + 3 * (size_t)(edge_ok ? jx : 0) * num_pts;
  const long long* __restrict__ idx_e = reinterpret_cast<const long long*>(idx) + eb;
#ifndef M3R_PACK_UNROLL
#define M3R_PACK_UNROLL 2
#endif
  constexpr int kPackUnroll = M3R_PACK_UNROLL;
#pragma unroll kPackUnroll
  for (int p = 0; p < kPasses; p++) {
    const int lk = p * 1024 + warp * 128 + lane * 4;  // pixel offset inside the tile
    const int k = tile * kTilePix + lk;
    float q[4] = {0.f, 0.f, 0.f, 0.f};
    unsigned vm = 0;
    long long ind[4] = {-1, -1, -1, -1};
    if (edge_ok && k < num_pts) {
      if (vec_ok) {  // k + 3 < num_pts because num_pts % 4 == 0
        const float4 qv = __ldg(reinterpret_cast<const float4*>(Q + eb + k));  // re-read in phase 2: keep it in L2
        q[0] = qv.x; q[1] = qv.y; q[2] = qv.z; q[3] = qv.w;
        vm = __ldcs(reinterpret_cast<const unsigned*>(valid + eb + k));
      } else {
#pragma unroll
        for (int j = 0; j < 4; j++)
          if (k + j < num_pts) {
            q[j] = __ldg(Q + eb + k + j);
            vm |= (unsigned)(valid[eb + k + j] != 0) << (8 * j);
          }
      }
      bool pass[4];
#pragma unroll
      for (int j = 0; j < 4; j++) pass[j] = ((vm >> (8 * j)) & 0xffu) != 0 && q[j] > Q_thresh;
      if (vec_ok) {
        if (pass[0] | pass[1]) {
          const longlong2 v = __ldcs(reinterpret_cast<const longlong2*>(idx_e + k));
          ind[0] = v.x; ind[1] = v.y;
        }
        if (pass[2] | pass[3]) {
          const longlong2 v = __ldcs(reinterpret_cast<const longlong2*>(idx_e + k + 2));
          ind[2] = v.x; ind[3] = v.y;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 4; j++)
          if (pass[j]) ind[j] = idx_e[k + j];
      }
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (!pass[j] || ind[j] < 0 || ind[j] >= num_pts) ind[j] = -1;  // corrupt indices never reach a gather
    }
    int id[4];
#pragma unroll
    for (int j = 0; j < 4; j++) id[j] = (int)ind[j];
    if (cgate) {
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (id[j] >= 0 && !(__ldg(Ci + id[j]) > C_thresh && __ldg(Cj + k + j) > C_thresh)) id[j] = -1;
    }
    if (zgate) {
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (id[j] >= 0 && !(__ldg(Xi + 3 * (size_t)id[j] + 2) > z_eps)) id[j] = -1;
    }
    *reinterpret_cast<int4*>(&s_idx[lk]) = make_int4(id[0], id[1], id[2], id[3]);
    const int c = (id[0] >= 0) + (id[1] >= 0) + (id[2] >= 0) + (id[3] >= 0);
    const int tot = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0) s_cnt[p * 8 + warp] = tot;
  }
  __syncthreads();
  if (warp == 0) {
    // exclusive scan of the 64 (pass, warp) counts in pixel order, two per lane
    const int c0 = s_cnt[2 * lane], c1 = s_cnt[2 * lane + 1];
    int incl = c0 + c1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    const int run = __shfl_sync(0xffffffffu, incl, 31);
    s_cnt[2 * lane] = incl - c0 - c1;
    s_cnt[2 * lane + 1] = incl - c1;
    if (lane == 0) st_release(&tstate[1 + work], run + 1);  // own count, visible before anyone needs it
    int before = 0;
    for (int t = lane; t < tile; t += 32) {  // counts of the earlier tiles of this edge
      int v;
      while ((v = ld_acquire(&tstate[1 + work - tile + t])) == 0) __nanosleep(32);
      before += v - 1;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
    if (lane == 0) {
      s_base = before;
      if (tile == tiles - 1) ecount[e] = before + run;
    }
  }
  __syncthreads();
  // ---- phase 2a: compact (match index, pixel) of the tile's survivors in shared memory, in pixel order.  Every thread
  //      first takes its 32 entries into registers; after the barrier the compacted list overwrites s_idx in place.
  const unsigned lt = (1u << lane) - 1u;
  const int base = s_base;
  int4 idr[kPasses];
#pragma unroll
  for (int p = 0; p < kPasses; p++) idr[p] = *reinterpret_cast<const int4*>(&s_idx[p * 1024 + warp * 128 + lane * 4]);
  __syncthreads();
  int total = 0;
#pragma unroll
  for (int p = 0; p < kPasses; p++) {
    const int lk = p * 1024 + warp * 128 + lane * 4;
    const int4 id = idr[p];
    const unsigned b0 = __ballot_sync(0xffffffffu, id.x >= 0), b1 = __ballot_sync(0xffffffffu, id.y >= 0);
    const unsigned b2 = __ballot_sync(0xffffffffu, id.z >= 0), b3 = __ballot_sync(0xffffffffu, id.w >= 0);
    int pos = s_cnt[p * 8 + warp] + __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
    if (id.x >= 0) { s_idx[pos] = id.x; s_lk[pos] = (unsigned short)(lk);     pos++; }
    if (id.y >= 0) { s_idx[pos] = id.y; s_lk[pos] = (unsigned short)(lk + 1); pos++; }
    if (id.z >= 0) { s_idx[pos] = id.z; s_lk[pos] = (unsigned short)(lk + 2); pos++; }
    if (id.w >= 0) { s_idx[pos] = id.w; s_lk[pos] = (unsigned short)(lk + 3); pos++; }
    if (p == kPasses - 1 && warp == 7) total = s_cnt[p * 8 + warp] + __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
  }
  if (warp == 7 && lane == 0) s_work = total;
  __syncthreads();
  total = s_work;
  // ---- phase 2b: one thread per surviving record: the gathers of consecutive records are independent (memory-level
  //      parallelism), the two 16-byte stores of a warp are contiguous
  for (int r = threadIdx.x; r < total; r += 256) {
    const int ind = s_idx[r];
    const int k = tile * kTilePix + (int)s_lk[r];
    const float* xj = Xj + 3 * (size_t)k;
    const float q = __ldg(Q + eb + k);
    const size_t pos = eb + (size_t)base + r;
    __stcs(&rec[pos].a, make_rec_a(kind, Xi, ind, width));
    __stcs(&rec[pos].b, make_float4(__ldg(xj), __ldg(xj + 1), __ldg(xj + 2), sqrtf(q)));
  }
}

// eoff[e] = sum of ecount[0..e) (long long: E*N can exceed 2^31).  Single CTA; E is a few thousand at most.
__global__ void __launch_bounds__(1024) edge_offsets_kernel(const int* __restrict__ ecount, long long* __restrict__ eoff,
                                                            int num_edges) {
  __shared__ long long s_part[1024];
  const int per = (num_edges + 1023) / 1024;
  const int b = threadIdx.x * per;
  long long sum = 0;
  for (int i = 0; i < per; i++)
    if (b + i < num_edges) sum += ecount[b + i];
  s_part[threadIdx.x] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long run = 0;
    for (int i = 0; i < 1024; i++) {
      const long long c = s_part[i];
      s_part[i] = run;
      run += c;
    }
    eoff[num_edges] = run;
  }
  __syncthreads();
  long long run = s_part[threadIdx.x];
  for (int i = 0; i < per; i++)
    if (b + i < num_edges) {
      eoff[b + i] = run;
      run += ecount[b + i];
    }
}

// ------------------------------------------------------------------------------------------------ robust weights
// All per-record arithmetic below runs on PAIRS of records (V2, f32x2.cuh): one thread carries record A in the low and
// record B in the high half of every register pair, so each FADD2/FMUL2/FFMA2 serves two records.
//
// Huber (gn_kernels.cu:172-175): w = |r| < k ? 1 : k/|r|  ==  min(1, k * rcp(|r|)), branch-free; r = 0 gives
// min(1, inf) = 1.  The reference compares against the double literal 1.345 and divides in double; rcp.approx is within
// 1 ulp of that quotient and the two forms agree except within 1 ulp of the switch point, where both give w = 1 - O(ulp)
// (documented deviation).  Tukey (nonlinear_optimizer.py:36-42): (1 - (r/t)^2)^2 inside |r| < t, else 0
// == max(1 - (r/t)^2, 0)^2.
template <int ROBUST>
__device__ __forceinline__ V2 robust_weight2(V2 r) {
  if (ROBUST == M3R_ROBUST_HUBER) {
    return min2(mul2(rcp_abs2(r), bc(1.345f)), 1.0f);
  } else {
    const V2 t = mul2(r, bc(1.0f / 4.6851f));
    const V2 m = max2(fma2(mul2(t, bc(-1.0f)), t, bc(1.0f)), 0.0f);
    return mul2(m, m);
  }
}

// acc[0..27]: lower triangle of S (row a, col b<=a at a(a+1)/2+b); acc[28..34]: g.
// MASK: bit a set <=> J[a] may be non-zero for this residual row (compile-time sparsity of the local Jacobian).
template <int MASK>
__device__ __forceinline__ void accum_row(V2 (&acc)[kAcc], const V2 (&J)[7], V2 w, V2 e) {
  V2 wJ[7];
#pragma unroll
  for (int a = 0; a < 7; a++)
    if ((MASK >> a) & 1) wJ[a] = mul2(w, J[a]);
  const V2 we = mul2(w, e);
#pragma unroll
  for (int a = 0; a < 7; a++) {
    if (!((MASK >> a) & 1)) continue;
#pragma unroll
    for (int b = 0; b <= a; b++) {
      if (!((MASK >> b) & 1)) continue;
      fma2_acc(acc[a * (a + 1) / 2 + b], wJ[a], J[b]);
    }
    fma2_acc(acc[28 + a], we, J[a]);
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

__device__ __forceinline__ float act_row(const float* R, float t, const float4& x) {
  return fmaf(R[0], x.x, fmaf(R[1], x.y, fmaf(R[2], x.z, t)));
}

// Two records of one edge: residuals, weights, local Jacobian rows, accumulation.  A padding lane carries sqrt(q) = 0
// and a copy of a real record: every weight is then exactly zero and every term finite.
template <int KIND, int ROBUST>
__device__ __forceinline__ void pair_accumulate(V2 (&acc)[kAcc], const EdgeConst& c, const PixParams& pp,
                                                const float4 a0, const float4 a1, const float4 b0, const float4 b1,
                                                V2 sq) {
  // P = s_ij R_ij X_j + t_ij: the edge constants are scalars, one FFMA per record and term
  const V2 Px = pk(act_row(c.R + 0, c.t[0], b0), act_row(c.R + 0, c.t[0], b1));
  const V2 Py = pk(act_row(c.R + 3, c.t[1], b0), act_row(c.R + 3, c.t[1], b1));
  V2 Pz = pk(act_row(c.R + 6, c.t[2], b0), act_row(c.R + 6, c.t[2], b1));
  V2 J[7];
  if (KIND == M3R_GN_RAYS) {
    const V2 n2j = fma2(Px, Px, fma2(Py, Py, mul2(Pz, Pz)));
    const V2 invj = rsqrt2(n2j);
    const V2 n1j = mul2(n2j, invj);
    const V2 rx = mul2(Px, invj), ry = mul2(Py, invj), rz = mul2(Pz, invj);
    const V2 e0 = sub2(rx, pk(a0.x, a1.x)), e1 = sub2(ry, pk(a0.y, a1.y)), e2 = sub2(rz, pk(a0.z, a1.z));
    const V2 e3 = sub2(n1j, pk(a0.w, a1.w));  // |P| - |X_i|
    const V2 sw_ray = mul2(sq, bc(pp.sa_inv)), sw_dist = mul2(sq, bc(pp.sb_inv));
    const V2 wc_ray = mul2(sw_ray, sw_ray), wc_dist = mul2(sw_dist, sw_dist);
    const V2 w0 = mul2(robust_weight2<ROBUST>(mul2(sw_ray, e0)), wc_ray);
    const V2 w1 = mul2(robust_weight2<ROBUST>(mul2(sw_ray, e1)), wc_ray);
    const V2 w2 = mul2(robust_weight2<ROBUST>(mul2(sw_ray, e2)), wc_ray);
    const V2 w3 = mul2(robust_weight2<ROBUST>(mul2(sw_dist, e3)), wc_dist);
    // d r / d P = (I - r r^T) / |P|   (gn_kernels.cu:982-988), with s = r/|P| and the negated ray n = -r
    const V2 neg1 = bc(-1.0f);
    const V2 nx = mul2(rx, neg1), ny = mul2(ry, neg1), nz = mul2(rz, neg1);
    const V2 sx = mul2(rx, invj), sy = mul2(ry, invj), sz = mul2(rz, invj);
    const V2 dxx = fma2(sx, nx, invj), dyy = fma2(sy, ny, invj), dzz = fma2(sz, nz, invj);
    const V2 dxy = mul2(sx, ny), dxz = mul2(sx, nz), dyz = mul2(sy, nz);
    J[0] = dxx; J[1] = dxy; J[2] = dxz; J[3] = dxx; J[4] = rz; J[5] = ny; J[6] = dxx;
    accum_row<0b0110111>(acc, J, w0, e0);
    J[0] = dxy; J[1] = dyy; J[2] = dyz; J[3] = nz; J[5] = rx;
    accum_row<0b0101111>(acc, J, w1, e1);
    J[0] = dxz; J[1] = dyz; J[2] = dzz; J[3] = ry; J[4] = nx;
    accum_row<0b0011111>(acc, J, w2, e2);
    J[0] = rx; J[1] = ry; J[2] = rz; J[6] = n1j;
    accum_row<0b1000111>(acc, J, w3, e3);
  } else if (KIND == M3R_GN_POINTS) {
    const V2 e0 = sub2(Px, pk(a0.x, a1.x)), e1 = sub2(Py, pk(a0.y, a1.y)), e2 = sub2(Pz, pk(a0.z, a1.z));
    const V2 sw = mul2(sq, bc(pp.sa_inv));
    const V2 wc = mul2(sw, sw);
    const V2 w0 = mul2(robust_weight2<ROBUST>(mul2(sw, e0)), wc);
    const V2 w1 = mul2(robust_weight2<ROBUST>(mul2(sw, e1)), wc);
    const V2 w2 = mul2(robust_weight2<ROBUST>(mul2(sw, e2)), wc);
    const V2 one = bc(1.0f), neg1 = bc(-1.0f);
    const V2 nPx = mul2(Px, neg1), nPy = mul2(Py, neg1), nPz = mul2(Pz, neg1);
    J[0] = one; J[1] = one; J[2] = one; J[3] = one; J[4] = Pz; J[5] = nPy; J[6] = Px;
    accum_row<0b1110001>(acc, J, w0, e0);
    J[3] = nPz; J[5] = Px; J[6] = Py;
    accum_row<0b1101010>(acc, J, w1, e1);
    J[3] = Py; J[4] = nPx; J[6] = Pz;
    accum_row<0b1011100>(acc, J, w2, e2);
  } else {  // M3R_GN_CALIB, gn_kernels.cu:1360-1479
    // per-record validity that depends on the poses (projected depth > eps, projection inside the bordered image): an
    // invalid record keeps a finite stand-in depth and gets sq = 0, i.e. weight exactly zero as in the reference
    // (:1399-1404); z_i > eps was gated by the pre-pass
    const bool zA = lo(Pz) > pp.z_eps, zB = hi(Pz) > pp.z_eps;
    Pz = pk(zA ? lo(Pz) : 1.0f, zB ? hi(Pz) : 1.0f);
    const V2 zj_inv = pk(1.0f / lo(Pz), 1.0f / hi(Pz));
    const V2 x = mul2(Px, zj_inv), y = mul2(Py, zj_inv);
    const V2 fx = bc(c.fx), fy = bc(c.fy);
    const V2 u = fma2(fx, x, bc(c.cx)), v = fma2(fy, y, bc(c.cy));
    const float ulo = (float)pp.pixel_border, uhi = (float)(pp.width - 1 - pp.pixel_border),
                vhi = (float)(pp.height - 1 - pp.pixel_border);
    const bool okA = zA && lo(u) > ulo && lo(u) < uhi && lo(v) > ulo && lo(v) < vhi;
    const bool okB = zB && hi(u) > ulo && hi(u) < uhi && hi(v) > ulo && hi(v) < vhi;
    sq = pk(okA ? lo(sq) : 0.0f, okB ? hi(sq) : 0.0f);
    const V2 e0 = sub2(u, pk(a0.x, a1.x)), e1 = sub2(v, pk(a0.y, a1.y));
    const V2 e2 = pk(logf(lo(Pz)) - a0.z, logf(hi(Pz)) - a1.z);
    const V2 sw_pix = mul2(sq, bc(pp.sa_inv)), sw_dep = mul2(sq, bc(pp.sb_inv));
    const V2 wc_pix = mul2(sw_pix, sw_pix), wc_dep = mul2(sw_dep, sw_dep);
    const V2 w0 = mul2(robust_weight2<ROBUST>(mul2(sw_pix, e0)), wc_pix);
    const V2 w1 = mul2(robust_weight2<ROBUST>(mul2(sw_pix, e1)), wc_pix);
    const V2 w2 = mul2(robust_weight2<ROBUST>(mul2(sw_dep, e2)), wc_dep);
    const V2 one = bc(1.0f), neg1 = bc(-1.0f);
    const V2 fxz = mul2(fx, zj_inv), fyz = mul2(fy, zj_inv);
    const V2 nx = mul2(x, neg1), ny = mul2(y, neg1);
    J[0] = fxz; J[1] = one; J[2] = mul2(fxz, nx); J[3] = mul2(mul2(fx, nx), y); J[4] = mul2(fx, fma2(x, x, one));
    J[5] = mul2(fx, ny); J[6] = one;
    accum_row<0b0111101>(acc, J, w0, e0);
    J[1] = fyz; J[2] = mul2(fyz, ny); J[3] = mul2(fy, fma2(ny, y, neg1)); J[4] = mul2(mul2(fy, x), y);
    J[5] = mul2(fy, x);
    accum_row<0b0111110>(acc, J, w1, e1);
    J[2] = zj_inv; J[3] = y; J[4] = nx; J[6] = one;
    accum_row<0b1011100>(acc, J, w2, e2);
  }
}

// ------------------------------------------------------------------------------------------------ Hessian kernel
// Persistent, stream-K style: the V = sum_e ecount[e] compacted records of all edges form one flat list; CTA c of the
// grid (kHessCtasPerSm CTAs per SM) owns the records [c*chunk, (c+1)*chunk) and walks the (at most a few) edge
// segments inside its range.  Per segment: edge constants once; the records arrive in shared memory as tiles of
// 2*kHessThreads through a kHessStages-deep ring of TMA bulk copies (cp.async.bulk + mbarrier: one elected thread arms
// the barrier and issues two copies per tile, nothing else is spent on the transfer); every thread takes two records of
// the tile (packed f32x2 arithmetic); 35 accumulators are reduced once per segment: transposing butterfly across lanes
// (31 shuffles for 32 values), shared memory across the warps.  The partial sums of segment (e, c) go to slot e + c,
// which is unique because consecutive segments advance e, c or both; finalize_edge sums an edge's slots in fixed order.
// History (profiles/): v1 one CTA per 2048-px tile, latency-bound at 25 % occupancy; v2 8192-px tiles + pipelined
// loads; v3 gated-out pixels dropped by the pre-pass, persistent CTAs; v4 gathered points through a cp.async ring
// (issue-bound, 236 instructions per record); v6 packed f32x2 arithmetic, MUFU-only reciprocals, branch-free Huber (143
// per record) -- no faster: the two 16-byte gathers per record saturated the L1TEX wavefront rate (~100 G records/s at 32
// and at 256 keyframes alike, profiles/r2_hessian_v6_ncu_full.txt); v7 (this): the gathers moved into the once-per-solve
// pre-pass, the kernel streams 32 self-contained bytes per record through one CTA-wide ring (83 us at configs[2], 1006 us
// at 256 keyframes); v8 (default, M3R_HESS_WARPRING=1): warp-private rings -- every warp streams its own quarter of a
// segment and refills a stage itself, so no warp waits for the slowest one before copies are issued: 79.6 / 955 us with 3
// stages (4: 82.1 / 966, 6: 80.0 / 959, 8: 81.1 / 968; profiles/r2_hessian_v8_variants.txt).  Also tried on v7 and dropped:
// 4 CTAs per SM under a 128-register cap (88 B of spills: 91 / 1141 us), 96-thread CTAs x 5 (118 / 1514 us).
#ifndef M3R_HESS_EVICT_FIRST
#define M3R_HESS_EVICT_FIRST 1  // record loads carry an L2 evict-first policy (tma.cuh)
#endif
#ifndef M3R_HESS_WARPRING
#define M3R_HESS_WARPRING 1
#endif
#ifndef M3R_HESS_STAGES
#define M3R_HESS_STAGES (M3R_HESS_WARPRING ? 3 : 4)
#endif
constexpr int kHessStages = M3R_HESS_STAGES;
constexpr int kHessTile = 2 * kHessThreads;  // records per ring stage
#if M3R_HESS_WARPRING
// Warp-private rings: every warp streams its own quarter of a segment through its own stages and refills a stage itself
// as soon as its 32 lanes have taken their records out -- no cross-warp "empty" barrier, so a slow warp never holds back
// the copies of the others.
constexpr int kHessWarps = kHessThreads / 32;
constexpr int kWarpTile = 64;  // records per warp and stage (two per lane)
struct HessRing {
  GnRecord r[kHessWarps][kHessStages][kWarpTile];
  unsigned long long full[kHessWarps][kHessStages];
};
#else
struct HessRing {
  GnRecord r[kHessStages][kHessTile];
  unsigned long long full[kHessStages];   // tile landed (TMA transaction bytes)
  unsigned long long empty[kHessStages];  // every warp has taken its records out of the stage
};
#endif

constexpr int kHessMaxSeg = 8;  // edge segments planned at once by a CTA (more are handled in further rounds)
struct HessPlan {
  EdgeConst c[kHessMaxSeg];
  int e[kHessMaxSeg], k0[kHessMaxSeg], k1[kHessMaxSeg];
  int nseg, e_next;
  long long p_next;
};

template <int KIND, int ROBUST>
__global__ void __launch_bounds__(kHessThreads, kHessCtasPerSm) hessian_kernel(const float* __restrict__ Twc,
                                                                  const float* __restrict__ K,
                                                                  const GnRecord* __restrict__ rec,
                                                                  const long long* __restrict__ eoff,
                                                                  const int* __restrict__ ie, const int* __restrict__ je,
                                                                  float* __restrict__ partial,
                                                                  const int* __restrict__ state, int num_pts,
                                                                  int num_edges, PixParams pp) {
  if (state[kDone]) return;
  const long long V = eoff[num_edges];
  const long long chunk = hess_chunk(V, gridDim.x);
  long long p = (long long)blockIdx.x * chunk;
  const long long p_end = min(V, p + chunk);
  if (p >= p_end) return;
  __shared__ HessPlan plan;
  __shared__ float s_red[2][kHessThreads / 32][kAccStride];
  extern __shared__ __align__(128) unsigned char s_ring_raw[];
  HessRing& ring = *reinterpret_cast<HessRing*>(s_ring_raw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#if M3R_HESS_WARPRING
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kHessStages; s++) mbar_init(&ring.full[warp][s], 1);
  }
#else
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kHessStages; s++) {
      mbar_init(&ring.full[s], 1);
      mbar_init(&ring.empty[s], kHessThreads / 32);
    }
  }
#endif
  // tiles issued / consumed so far by this CTA: stage = count % kHessStages, phase parity from count / kHessStages
  unsigned issued = 0, consumed = 0, nred = 0;
  int e_cur = -1;
  while (p < p_end) {
    // ---- plan the next (up to kHessMaxSeg) edge segments of the CTA's record range: thread 0 finds the edge that owns
    //      record p (binary search once, then a forward walk over empty edges) and the segment bounds ...
    //      (Tried, r2: a 32-way search by warp 0 with the next 32 offsets fetched in one go, and the first copies issued
    //      before the edge constants are built: 78.8 against 78.1 us at configs[2] -- the offsets sit in L1/L2 and the
    //      plan is not where the ~13 us of fixed cost per launch go.)
    if (threadIdx.x == 0) {
      int en;
      if (e_cur < 0) {
        int lo_e = 0, hi_e = num_edges - 1;  // last edge with eoff[edge] <= p
        while (lo_e < hi_e) {
          const int mid = (lo_e + hi_e + 1) >> 1;
          if (eoff[mid] <= p) lo_e = mid; else hi_e = mid - 1;
        }
        en = lo_e;
      } else {
        en = e_cur;
      }
      int n = 0;
      long long q = p;
      while (q < p_end && n < kHessMaxSeg) {
        long long nxt;
        while ((nxt = eoff[en + 1]) <= q) en++;
        const long long b0 = eoff[en];
        const long long seg_end = min(p_end, nxt);
        plan.e[n] = en;
        plan.k0[n] = (int)(q - b0);
        plan.k1[n] = (int)(seg_end - b0);
        n++;
        q = seg_end;
        en++;
      }
      plan.nseg = n;
      plan.p_next = q;
      plan.e_next = en;
    }
    __syncthreads();  // also orders the mbarrier initialisation before their first use
    // ... and one thread per segment builds its edge constants
    const int nseg = plan.nseg;
    if ((int)threadIdx.x < nseg)
      edge_setup(Twc, ie[plan.e[threadIdx.x]], je[plan.e[threadIdx.x]], KIND == M3R_GN_CALIB ? K : nullptr,
                 &plan.c[threadIdx.x]);
    __syncthreads();

#if M3R_HESS_WARPRING
    // This warp's slice of planned segment sidx: a quarter of the segment, rounded up to whole warp tiles.
    auto warp_slice = [&](int sidx, int& ws, int& we) {
      const int k0 = plan.k0[sidx], k1 = plan.k1[sidx];
      const int q = ((k1 - k0 + kHessWarps * kWarpTile - 1) / (kHessWarps * kWarpTile)) * kWarpTile;
      ws = min(k1, k0 + warp * q);
      we = min(k1, ws + q);
    };
    // Producer cursor of the warp (kept uniform over the lanes; lane 0 issues): the warp tiles of all planned segments
    // form one sequence, so the copies run ahead ACROSS segment boundaries.
    int pseg = 0, pkb = 0, pend = 0;
    warp_slice(0, pkb, pend);
#if M3R_HESS_EVICT_FIRST
    const unsigned long long l2pol = l2_policy_evict_first();
#endif
    auto issue_next = [&]() {
      while (pseg < nseg && pkb >= pend) {  // skip segments in which this warp has no records (left)
        pseg++;
        if (pseg < nseg) warp_slice(pseg, pkb, pend);
      }
      if (pseg >= nseg) return;
      const unsigned st = issued % kHessStages;
      if (lane == 0) {
        const unsigned bytes = (unsigned)min(kWarpTile, pend - pkb) * (unsigned)sizeof(GnRecord);
        mbar_arrive_expect_tx(&ring.full[warp][st], bytes);
#if M3R_HESS_EVICT_FIRST
        bulk_g2s_hint(&ring.r[warp][st][0], rec + (size_t)plan.e[pseg] * num_pts + pkb, bytes, &ring.full[warp][st], l2pol);
#else
        bulk_g2s(&ring.r[warp][st][0], rec + (size_t)plan.e[pseg] * num_pts + pkb, bytes, &ring.full[warp][st]);
#endif
      }
      issued++;
      pkb += kWarpTile;
    };
#pragma unroll
    for (int i = 0; i < kHessStages; i++) issue_next();

    for (int sg = 0; sg < nseg; sg++) {
      const EdgeConst c = plan.c[sg];
      const int e = plan.e[sg];
      int ws, we;
      warp_slice(sg, ws, we);
      V2 acc2[kAcc];
#pragma unroll
      for (int i = 0; i < kAcc; i++) acc2[i] = bc(0.0f);
#pragma unroll 2
      for (int kb = ws; kb < we; kb += kWarpTile) {
        const unsigned st = consumed % kHessStages;
        mbar_wait(&ring.full[warp][st], (consumed / kHessStages) & 1);
        // The tile holds cnt >= 1 records; a lane without a record of its own re-reads the last one with sqrt(q) = 0
        // (weight exactly zero, all terms finite): no branch around the arithmetic and its 70 accumulator registers.
        const int cnt = min(kWarpTile, we - kb);
        const int i0 = min(lane, cnt - 1), i1 = min(lane + 32, cnt - 1);
        const float4 a0 = ring.r[warp][st][i0].a, b0 = ring.r[warp][st][i0].b;
        const float4 a1 = ring.r[warp][st][i1].a, b1 = ring.r[warp][st][i1].b;
        __syncwarp();  // every lane has its records in registers: the stage is free
        if (lane == 0) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // reads before the copy engine's write
        consumed++;
        issue_next();  // refills exactly the stage just freed (issued == consumed + kHessStages - 1 before the call)
        const float sq0 = lane < cnt ? b0.w : 0.0f;
        const float sq1 = lane + 32 < cnt ? b1.w : 0.0f;
        pair_accumulate<KIND, ROBUST>(acc2, c, pp, a0, a1, b0, b1, pk(sq0, sq1));
      }
#else
    // Producer cursor (thread 0 only): the tiles of all planned segments form one sequence, so the ring keeps
    // kHessStages - 1 tiles in flight ACROSS segment boundaries.  (Tried, r2: the role rotating over the warps, 83 ->
    // 117 us; deeper rings, 4 -> 6 -> 8 stages: 83 -> 94 -> 100 us; an L2 prefetch 8 / 16 tiles ahead: 99 / 109 us; two
    // tiles per loop trip: 88 us; two float4 streams instead of one 32-byte record stream: the same 83 us.)
    int pseg = 0, pkb = plan.k0[0];
    auto issue_next = [&]() {
      if (pseg >= nseg) return;
      const unsigned st = issued % kHessStages, round = issued / kHessStages;
      if (round > 0) mbar_wait(&ring.empty[st], (round - 1) & 1);  // the stage's previous tile has been taken out
      const int e = plan.e[pseg], k1 = plan.k1[pseg];
      const unsigned bytes = (unsigned)min(kHessTile, k1 - pkb) * (unsigned)sizeof(GnRecord);
#ifndef M3R_HESS_NOLOAD
      mbar_arrive_expect_tx(&ring.full[st], bytes);
      bulk_g2s(&ring.r[st][0], rec + (size_t)e * num_pts + pkb, bytes, &ring.full[st]);
#endif
      issued++;
      pkb += kHessTile;
      if (pkb >= k1) {
        pseg++;
        if (pseg < nseg) pkb = plan.k0[pseg];
      }
    };
    if (threadIdx.x == 0) {
#pragma unroll
      for (int i = 0; i < kHessStages - 1; i++) issue_next();
    }

    for (int sg = 0; sg < nseg; sg++) {
      const EdgeConst c = plan.c[sg];
      const int e = plan.e[sg], k0 = plan.k0[sg], k1 = plan.k1[sg];
      const int ntiles = (k1 - k0 + kHessTile - 1) / kHessTile;
      V2 acc2[kAcc];
#pragma unroll
      for (int i = 0; i < kAcc; i++) acc2[i] = bc(0.0f);
#pragma unroll 2
      for (int t = 0; t < ntiles; t++) {
        if (threadIdx.x == 0) issue_next();
        const unsigned st = consumed % kHessStages;
#ifndef M3R_HESS_NOLOAD  // experiment: compute side only (no copies issued, nothing to wait for)
        mbar_wait(&ring.full[st], (consumed / kHessStages) & 1);
#endif
        // The tile holds cnt >= 1 records.  Every thread always computes: a thread without a record of its own re-reads
        // the tile's last record with sqrt(q) = 0 (weight exactly zero, all terms finite) -- no branch around the
        // arithmetic, whose 70 accumulator registers would otherwise have to be merged behind it.
        const int cnt = min(kHessTile, k1 - (k0 + t * kHessTile));
        const int i0 = min((int)threadIdx.x, cnt - 1), i1 = min((int)threadIdx.x + kHessThreads, cnt - 1);
        const float4 a0 = ring.r[st][i0].a, b0 = ring.r[st][i0].b;
        const float4 a1 = ring.r[st][i1].a, b1 = ring.r[st][i1].b;
        __syncwarp();  // every lane has its records in registers; one arrival per warp frees the stage
        if (lane == 0) mbar_arrive(&ring.empty[st]);
        consumed++;
        const float sq0 = (int)threadIdx.x < cnt ? b0.w : 0.0f;
        const float sq1 = (int)threadIdx.x + kHessThreads < cnt ? b1.w : 0.0f;
#ifdef M3R_HESS_NOCOMPUTE  // experiment: memory side only
        acc2[0] = add2(acc2[0], pk(a0.x + b0.x + sq0, a1.x + b1.x + sq1));
#else
        pair_accumulate<KIND, ROBUST>(acc2, c, pp, a0, a1, b0, b1, pk(sq0, sq1));
#endif
      }
#endif
      float acc[kAcc];
#pragma unroll
      for (int i = 0; i < kAcc; i++) acc[i] = lo(acc2[i]) + hi(acc2[i]);
      // ---- block reduction: transposing butterfly for acc[0..31] (lane l ends up with the warp total of acc[l]),
      //      plain xor-tree for the remaining kAcc-32 values, then the warps through shared memory (two buffers in
      //      turn: the next segment's writers can only be one barrier ahead of this segment's readers)
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
      float(*red)[kAccStride] = s_red[nred & 1];
      nred++;
      red[warp][lane] = acc[0];
#pragma unroll
      for (int i = 32; i < kAcc; i++) {
        float v = acc[i];
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        if (lane == 0) red[warp][i] = v;
      }
      __syncthreads();
      if (threadIdx.x < kAcc) {
        float v = 0.0f;
#pragma unroll
        for (int wv = 0; wv < kHessThreads / 32; wv++) v += red[wv][threadIdx.x];
        partial[(size_t)(e + (int)blockIdx.x) * kAccStride + threadIdx.x] = v;
      }
    }
    p = plan.p_next;
    e_cur = plan.e_next;
    __syncthreads();  // the plan is rewritten by the next round
  }
}

// ------------------------------------------------------------------------------------------------ per-edge finalize
// Diagnostics path only (m3r_gn_edge_blocks); the solver runs finalize_edge inside its fused kernel (gn_fused.cu).
__global__ void __launch_bounds__(128) finalize_edges_kernel(const float* Twc, const float* partial,
                                                             const long long* eoff, const int* ie, const int* je,
                                                             double* blocks, const int* state, int num_edges,
                                                             int hess_grid) {
  if (state[kDone]) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int e = blockIdx.x * 4 + warp;
  __shared__ FinalizeScratch sc[4];
  if (e >= num_edges) return;
  finalize_edge(e, lane, sc[warp], Twc, partial, eoff, ie, je, blocks, num_edges, hess_grid);
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
int launch_pack_points(const GnWs& ws, const float* Xs, const float* Cs, float C_thresh, cudaStream_t st) {
  (void)Xs;  // points are materialised per record by pack_edges; only the confidence scan is left of this pass
  const long long total = (long long)ws.L.num_kf * ws.L.num_pts;
  if (total == 0) return M3R_OK;
  long long blocks = ceil_div_ll(total, 256 * 8);
  if (blocks > 4 * sm_count()) blocks = 4 * sm_count();
  conf_scan_kernel<<<(unsigned)blocks, 256, 0, st>>>(Cs, total, C_thresh, ws.state() + kCFail);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int launch_pack_edges(const GnWs& ws, const m3r_gn_params& p, const float* Xs, const float* Cs, const int64_t* idx,
                      const uint8_t* valid, const float* Q, cudaStream_t st) {
  const int E = ws.L.num_edges, tiles = ws.L.tiles;
  M3R_CUDA(cudaMemsetAsync(ws.ecount(), 0, sizeof(int) * (size_t)(E + 1), st));
  if (E > 0 && tiles > 0) {
    M3R_CUDA(cudaMemsetAsync(ws.tstate(), 0, sizeof(int) * ((size_t)E * tiles + 2), st));
    static DynSmemOptIn attr;  // 48 KB dynamic + the static counters: above the default limit, opt in per device
    {
      const int rc = ensure_dyn_smem(pack_edges_kernel, 64 * 1024, attr);
      if (rc != M3R_OK) return rc;
    }
    pack_edges_kernel<<<(unsigned)((long long)E * tiles), 256, kPackSmemBytes, st>>>(
        Xs, Cs, idx, valid, Q, ws.ie(), ws.je(), ws.rec(), ws.ecount(), ws.tstate(), ws.state() + kCFail,
        ws.L.num_pts, tiles, p.C_thresh, p.Q_thresh, p.kind, p.width > 0 ? p.width : 1, p.z_eps);
    M3R_LAUNCH_CHECK();
  }
  edge_offsets_kernel<<<1, 1024, 0, st>>>(ws.ecount(), ws.eoff(), E);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int hessian_grid() {
  const int g = kHessCtasPerSm * sm_count();
  return g < kMaxHessCtas ? g : kMaxHessCtas;
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
  if (ws.L.num_edges == 0 || ws.L.num_pts == 0) return M3R_OK;
  const unsigned grid = (unsigned)hessian_grid();
  static DynSmemOptIn attr_t, attr_h;  // per instantiation (function-local statics of a template), per device
  {
    int rc = ensure_dyn_smem(hessian_kernel<KIND, M3R_ROBUST_TUKEY>, sizeof(HessRing), attr_t);
    if (rc == M3R_OK) rc = ensure_dyn_smem(hessian_kernel<KIND, M3R_ROBUST_HUBER>, sizeof(HessRing), attr_h);
    if (rc != M3R_OK) return rc;
  }
  if (p.robust == M3R_ROBUST_TUKEY)
    hessian_kernel<KIND, M3R_ROBUST_TUKEY><<<grid, kHessThreads, sizeof(HessRing), st>>>(
        Twc, ws.Kdev(), ws.rec(), ws.eoff(), ws.ie(), ws.je(), ws.partial(), ws.state(), ws.L.num_pts,
        ws.L.num_edges, pp);
  else
    hessian_kernel<KIND, M3R_ROBUST_HUBER><<<grid, kHessThreads, sizeof(HessRing), st>>>(
        Twc, ws.Kdev(), ws.rec(), ws.eoff(), ws.ie(), ws.je(), ws.partial(), ws.state(), ws.L.num_pts,
        ws.L.num_edges, pp);
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
  finalize_edges_kernel<<<ceil_div(ws.L.num_edges, 4), 128, 0, st>>>(Twc, ws.partial(), ws.eoff(), ws.ie(), ws.je(),
                                                                      ws.blocks(), ws.state(), ws.L.num_edges,
                                                                      hessian_grid());
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
