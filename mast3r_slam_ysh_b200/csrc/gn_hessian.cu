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
//   * Persistent stream-K decomposition over the flat list of compacted records (the reference runs one 256-thread
//     block per edge), two records per thread in packed f32x2 arithmetic, fp32 warp-shuffle + shared-memory tree per
//     segment, deterministic fp64 second stage across segments (no atomics).
#include "f32x2.cuh"
#include "gn_finalize.cuh"

namespace m3r {

// ------------------------------------------------------------------------------------------------ pre-pass
// Also answers "does any point of any keyframe fail the confidence gate c > C_thresh?" (state[kCFail]); when none does
// -- the usual case, C_conf = 0 and MASt3R confidences are >= 1 -- pack_edges skips its two confidence gathers per pixel.
__global__ void __launch_bounds__(256) pack_points_kernel(const float* __restrict__ Xs, const float* __restrict__ Cs,
                                                          float4* __restrict__ xc, long long total, float C_thresh,
                                                          int* __restrict__ cfail) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  bool fail = false;
  if (p < total) {
    const float c = __ldg(Cs + p);
    fail = !(c > C_thresh);  // NaN fails, as in the gate itself
    const float x = __ldg(Xs + 3 * p), y = __ldg(Xs + 3 * p + 1), z = __ldg(Xs + 3 * p + 2);
    // .w = -1/|X|: the Hessian kernel gets r_i = X_i/|X_i| and |X_i| without a rsqrt chain per record and iteration
    // (the confidence is only needed by the gates, which read Cs directly)
    xc[p] = make_float4(x, y, z, -1.0f / sqrtf(fmaf(x, x, fmaf(y, y, z * z))));
  }
  if (__any_sync(0xffffffffu, fail) && (threadIdx.x & 31) == 0) atomicOr(cfail, 1);
}

// Gate + compact.  One CTA per (edge, 8192-pixel tile); work items are handed out by a ticket so that the owners of
// the earlier tiles of an edge are guaranteed to be running (or finished) when a CTA looks back at them.
//   phase 1: every warp walks 1024 consecutive pixels (coalesced), evaluates the pose-independent gates and parks the
//            match index (or -1) in shared memory; survivors are counted with ballots.  The 8-byte index is only
//            fetched where valid_match and the Q gate pass, so whole sectors of it are never read outside the overlap
//            region of the two frames.
//   scan   : the CTA publishes its own count at once, then warp 0 sums the counts of the earlier tiles of the same edge
//            (decoupled look-back: nobody waits on a chain, only on each predecessor's phase 1).
//   phase 2: warps re-walk their pixels and write the survivors, in pixel order, to rec/sqv at the scanned offset.
__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(256) pack_edges_kernel(const float* __restrict__ Cs, const int64_t* __restrict__ idx,
                                                         const uint8_t* __restrict__ valid, const float* __restrict__ Q,
                                                         const int* __restrict__ ie, const int* __restrict__ je,
                                                         EdgeRec* __restrict__ rec, float* __restrict__ sqv,
                                                         int* __restrict__ ecount, int* __restrict__ tstate,
                                                         const int* __restrict__ cfail, int num_pts, int tiles,
                                                         float C_thresh, float Q_thresh) {
  // Every lane handles 4 consecutive pixels per pass (one LDG.32 of valid, one LDG.128 of Q, two predicated LDG.128 of
  // the 8-byte indices), 8 passes of 1024 pixels per tile: a quarter of the instructions of the one-pixel-per-lane
  // version, which was issue-bound at 52 % with 208 M warp instructions (profiles/r1_pack_v3a_ncu.txt).
  constexpr int kPasses = kTilePix / 1024;
  __shared__ __align__(16) int s_idx[kTilePix];
  __shared__ int s_work, s_base, s_cnt[kPasses * 8];
  if (threadIdx.x == 0) s_work = atomicAdd(&tstate[0], 1);
  __syncthreads();
  const int work = s_work;
  const int e = work / tiles, tile = work - e * tiles;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ix = ie[e], jx = je[e];
  const bool edge_ok = ix >= 0 && jx >= 0;
  const bool cgate = __ldg(cfail) != 0;  // some point fails c > C_thresh: evaluate the confidence gates per pixel
  const size_t eb = (size_t)e * num_pts;
  // vector path: every edge row of Q / idx starts 16-byte aligned and of valid 4-byte aligned
  const bool vec_ok = ((num_pts & 3) == 0) && ((reinterpret_cast<uintptr_t>(Q) & 15) == 0) &&
                      ((reinterpret_cast<uintptr_t>(idx) & 15) == 0) && ((reinterpret_cast<uintptr_t>(valid) & 3) == 0);
  const float* __restrict__ Ci = Cs + (size_t)(edge_ok ? ix : 0) * num_pts;
  const float* __restrict__ Cj = Cs + (size_t)(edge_ok ? jx : 0) * num_pts;
  const long long* __restrict__ idx_e = reinterpret_cast<const long long*>(idx) + eb;
#pragma unroll 2
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
        if (!pass[j] || ind[j] < 0 || ind[j] >= num_pts) ind[j] = -1;  // corrupt indices never reach the gather
    }
    int4 id;
    if (cgate) {
      float ci[4], cj[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        ci[j] = ind[j] >= 0 ? __ldg(Ci + ind[j]) : 0.0f;
        cj[j] = ind[j] >= 0 ? __ldg(Cj + k + j) : 0.0f;
      }
      id.x = (ind[0] >= 0 && ci[0] > C_thresh && cj[0] > C_thresh) ? (int)ind[0] : -1;
      id.y = (ind[1] >= 0 && ci[1] > C_thresh && cj[1] > C_thresh) ? (int)ind[1] : -1;
      id.z = (ind[2] >= 0 && ci[2] > C_thresh && cj[2] > C_thresh) ? (int)ind[2] : -1;
      id.w = (ind[3] >= 0 && ci[3] > C_thresh && cj[3] > C_thresh) ? (int)ind[3] : -1;
    } else {  // every confidence of every keyframe passes (pack_points checked them all)
      id.x = (int)ind[0];
      id.y = (int)ind[1];
      id.z = (int)ind[2];
      id.w = (int)ind[3];
    }
    *reinterpret_cast<int4*>(&s_idx[lk]) = id;
    const int c = (id.x >= 0) + (id.y >= 0) + (id.z >= 0) + (id.w >= 0);
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
  const unsigned lt = (1u << lane) - 1u;
  const int base = s_base;
#pragma unroll 2
  for (int p = 0; p < kPasses; p++) {
    const int lk = p * 1024 + warp * 128 + lane * 4;
    const int k = tile * kTilePix + lk;
    const int4 id = *reinterpret_cast<const int4*>(&s_idx[lk]);
    const unsigned b0 = __ballot_sync(0xffffffffu, id.x >= 0), b1 = __ballot_sync(0xffffffffu, id.y >= 0);
    const unsigned b2 = __ballot_sync(0xffffffffu, id.z >= 0), b3 = __ballot_sync(0xffffffffu, id.w >= 0);
    if ((b0 | b1 | b2 | b3) == 0) continue;
    int pos = base + s_cnt[p * 8 + warp] + __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
    {
      float q[4];
      if (vec_ok && k < num_pts) {
        const float4 qv = __ldcs(reinterpret_cast<const float4*>(Q + eb + k));
        q[0] = qv.x; q[1] = qv.y; q[2] = qv.z; q[3] = qv.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; j++) q[j] = (k + j < num_pts) ? Q[eb + k + j] : 0.0f;
      }
      if (id.x >= 0) { rec[eb + pos] = make_int2(id.x, k);     sqv[eb + pos] = sqrtf(q[0]); pos++; }
      if (id.y >= 0) { rec[eb + pos] = make_int2(id.y, k + 1); sqv[eb + pos] = sqrtf(q[1]); pos++; }
      if (id.z >= 0) { rec[eb + pos] = make_int2(id.z, k + 2); sqv[eb + pos] = sqrtf(q[2]); pos++; }
      if (id.w >= 0) { rec[eb + pos] = make_int2(id.w, k + 3); sqv[eb + pos] = sqrtf(q[3]); pos++; }
    }
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

// Two records of one edge: residuals, weights, local Jacobian rows, accumulation.  Record B may be a padding lane
// (sq.hi == 0, inputs copied from A): every weight is then exactly zero and every term finite.
// xi.w = -1/|X_i| (pack_points_kernel): r_i and |X_i| cost no rsqrt chain here.
template <int KIND, int ROBUST>
__device__ __forceinline__ void pair_accumulate(V2 (&acc)[kAcc], const EdgeConst& c, const PixParams& pp,
                                                const float4 xiA, const float4 xiB, const float4 xjA,
                                                const float4 xjB, int indA, int indB, V2 sq) {
  // P = s_ij R_ij X_j + t_ij: the edge constants are scalars, one FFMA per record and term
  const V2 Px = pk(act_row(c.R + 0, c.t[0], xjA), act_row(c.R + 0, c.t[0], xjB));
  const V2 Py = pk(act_row(c.R + 3, c.t[1], xjA), act_row(c.R + 3, c.t[1], xjB));
  V2 Pz = pk(act_row(c.R + 6, c.t[2], xjA), act_row(c.R + 6, c.t[2], xjB));
  V2 J[7];
  if (KIND == M3R_GN_RAYS) {
    const V2 n2j = fma2(Px, Px, fma2(Py, Py, mul2(Pz, Pz)));
    const V2 invj = rsqrt2(n2j);
    const V2 n1j = mul2(n2j, invj);
    const V2 rx = mul2(Px, invj), ry = mul2(Py, invj), rz = mul2(Pz, invj);
    const V2 wi = pk(xiA.w, xiB.w);  // -1/|X_i|
    const V2 e0 = fma2(pk(xiA.x, xiB.x), wi, rx), e1 = fma2(pk(xiA.y, xiB.y), wi, ry),
             e2 = fma2(pk(xiA.z, xiB.z), wi, rz);
    const V2 e3 = add2(n1j, rcp2(wi));  // |P| - |X_i|
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
    const V2 e0 = sub2(Px, pk(xiA.x, xiB.x)), e1 = sub2(Py, pk(xiA.y, xiB.y)), e2 = sub2(Pz, pk(xiA.z, xiB.z));
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
    // per-record validity (z > eps on both sides, projection inside the bordered image): an invalid record keeps finite
    // stand-in values and gets sq = 0, i.e. weight exactly zero as in the reference (:1399-1404)
    const bool zA = lo(Pz) > pp.z_eps && xiA.z > pp.z_eps, zB = hi(Pz) > pp.z_eps && xiB.z > pp.z_eps;
    Pz = pk(zA ? lo(Pz) : 1.0f, zB ? hi(Pz) : 1.0f);
    const float ziA = zA ? xiA.z : 1.0f, ziB = zB ? xiB.z : 1.0f;
    const V2 zj_inv = pk(1.0f / lo(Pz), 1.0f / hi(Pz));
    const V2 x = mul2(Px, zj_inv), y = mul2(Py, zj_inv);
    const V2 fx = bc(c.fx), fy = bc(c.fy);
    const V2 u = fma2(fx, x, bc(c.cx)), v = fma2(fy, y, bc(c.cy));
    const float ulo = (float)pp.pixel_border, uhi = (float)(pp.width - 1 - pp.pixel_border),
                vhi = (float)(pp.height - 1 - pp.pixel_border);
    const bool okA = zA && lo(u) > ulo && lo(u) < uhi && lo(v) > ulo && lo(v) < vhi;
    const bool okB = zB && hi(u) > ulo && hi(u) < uhi && hi(v) > ulo && hi(v) < vhi;
    sq = pk(okA ? lo(sq) : 0.0f, okB ? hi(sq) : 0.0f);
    const int vtA = indA / pp.width, vtB = indB / pp.width;
    const int utA = indA - vtA * pp.width, utB = indB - vtB * pp.width;
    const V2 e0 = sub2(u, pk((float)utA, (float)utB)), e1 = sub2(v, pk((float)vtA, (float)vtB));
    const V2 e2 = pk(logf(lo(Pz)) - logf(ziA), logf(hi(Pz)) - logf(ziB));
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
// grid (kHessCtasPerSm CTAs per SM) owns the records [c*chunk, (c+1)*chunk).  A CTA walks the (at most a few) edge
// segments inside its range; per segment it builds the edge constants once, streams the records through a software
// pipeline (records two batches ahead in registers, the gathered points one batch ahead in a cp.async shared-memory
// ring), and reduces its 35 accumulators once: transposing butterfly across lanes (31 shuffles for 32 values), shared
// memory across the warps.  The partial sums of segment (e, c) go to slot e + c, which is unique because consecutive
// segments advance e, c or both; finalize_edge sums an edge's slots in fixed order (no atomics).
// History (profiles/): v1 one CTA per 2048-px tile, latency-bound at 25 % occupancy; v2 8192-px tiles + pipelined
// loads; v3 drops gated-out pixels in the pre-pass and amortises the serial edge set-up over ~1e5 records per CTA; v4
// cp.async ring; v6 (this): two records per thread in packed f32x2 arithmetic, MUFU-only reciprocals, branch-free
// Huber, pre-normalised X_i -- ~236 -> ~105 issue slots per record.
struct RecBatch {
  int2 r[2];
  float sq[2];
};

__device__ __forceinline__ void load_recs(const EdgeRec* __restrict__ er, const float* __restrict__ es, int kb, int k1,
                                          RecBatch& rb) {
#pragma unroll
  for (int u = 0; u < 2; u++) {
    const int k = kb + u * kHessThreads;
    rb.r[u] = make_int2(-1, 0);
    rb.sq[u] = 0.0f;
    if (k < k1) {
      rb.r[u] = __ldcs(er + k);  // streamed once per iteration: evict-first in L2
      rb.sq[u] = __ldcs(es + k);
    }
  }
}

// The gathered points go global -> shared memory with cp.async (LDGSTS), one batch ahead of the arithmetic, into a
// two-slot per-thread ring: no registers are held by loads in flight.  A thread only ever touches its own slots, so
// there is no barrier in the loop.  (r1: ring depth 1 / 2 / 3 / 4 = 99.4 / 101.3 / 103.0 / 105.1 us at 32 kf.)
constexpr int kRingSlots = 2;
struct HessRing {
  float4 xi[kRingSlots][2][kHessThreads];
  float4 xj[kRingSlots][2][kHessThreads];
  float2 meta[kRingSlots][2][kHessThreads];  // (idx as int bits, sq)
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  // .cg (L2 only): the gathers have no L1 reuse (hit rate 10 %); with .ca the 256-keyframe case, whose 805 MB of
  // points do not fit L2, ran 1.60 ms instead of 1.32 ms (r1)
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

__device__ __forceinline__ void ring_issue(HessRing& ring, int slot, const float4* __restrict__ xci,
                                           const float4* __restrict__ xcj, const RecBatch& rb) {
#pragma unroll
  for (int u = 0; u < 2; u++) {
    ring.meta[slot][u][threadIdx.x] = make_float2(__int_as_float(rb.r[u].x), rb.sq[u]);
    if (rb.r[u].x >= 0) {
      cp_async16(&ring.xj[slot][u][threadIdx.x], xcj + rb.r[u].y);
      cp_async16(&ring.xi[slot][u][threadIdx.x], xci + rb.r[u].x);
    }
  }
  cp_async_commit();
}

template <int KIND, int ROBUST>
__global__ void __launch_bounds__(kHessThreads, kHessCtasPerSm) hessian_kernel(const float* __restrict__ Twc,
                                                                  const float* __restrict__ K,
                                                                  const float4* __restrict__ xc,
                                                                  const EdgeRec* __restrict__ rec,
                                                                  const float* __restrict__ sqv,
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
  __shared__ EdgeConst s_c;
  __shared__ int s_e;
  __shared__ float s_red[kHessThreads / 32][kAccStride];
  extern __shared__ __align__(16) unsigned char s_ring_raw[];
  HessRing& ring = *reinterpret_cast<HessRing*>(s_ring_raw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int e = -1;
  while (p < p_end) {
    // ---- next segment: thread 0 finds the edge that owns record p (binary search once, then a forward walk over
    //      empty edges) and builds the edge constants
    if (threadIdx.x == 0) {
      int en;
      if (e < 0) {
        int lo_e = 0, hi_e = num_edges - 1;  // last edge with eoff[edge] <= p
        while (lo_e < hi_e) {
          const int mid = (lo_e + hi_e + 1) >> 1;
          if (eoff[mid] <= p) lo_e = mid; else hi_e = mid - 1;
        }
        en = lo_e;
      } else {
        en = e + 1;
      }
      while (eoff[en + 1] <= p) en++;
      s_e = en;
      edge_setup(Twc, ie[en], je[en], KIND == M3R_GN_CALIB ? K : nullptr, &s_c);
    }
    __syncthreads();
    e = s_e;
    const EdgeConst c = s_c;
    const long long e0 = eoff[e];
    const long long seg_end = min(p_end, eoff[e + 1]);
    const int k1 = (int)(seg_end - e0);
    const float4* __restrict__ xci = xc + (size_t)ie[e] * num_pts;
    const float4* __restrict__ xcj = xc + (size_t)je[e] * num_pts;
    const EdgeRec* __restrict__ er = rec + (size_t)e * num_pts;
    const float* __restrict__ es = sqv + (size_t)e * num_pts;

    V2 acc2[kAcc];
#pragma unroll
    for (int i = 0; i < kAcc; i++) acc2[i] = bc(0.0f);
    {
      constexpr int kStep = 2 * kHessThreads;
      int kb = (int)(p - e0) + threadIdx.x;
      {  // prologue: gathers of batch 0 in flight, records of the next two batches in registers
        RecBatch r0;
        load_recs(er, es, kb, k1, r0);
        ring_issue(ring, 0, xci, xcj, r0);
      }
      RecBatch ra, rb;
      load_recs(er, es, kb + kStep, k1, ra);
      load_recs(er, es, kb + 2 * kStep, k1, rb);
      int slot = 0;
#pragma unroll 2
      for (; kb < k1; kb += kStep) {
        RecBatch rc;
        load_recs(er, es, kb + 3 * kStep, k1, rc);
        cp_async_wait<0>();  // this batch has landed
        const float2 mA = ring.meta[slot][0][threadIdx.x], mB = ring.meta[slot][1][threadIdx.x];
        const float4 xiA = ring.xi[slot][0][threadIdx.x], xjA = ring.xj[slot][0][threadIdx.x];
        float4 xiB = ring.xi[slot][1][threadIdx.x], xjB = ring.xj[slot][1][threadIdx.x];
        ring_issue(ring, slot ^ 1, xci, xcj, ra);
        // kb < k1 inside the loop, so record A always exists; records are dense, so B = A + kHessThreads can only be
        // missing at the end of the segment: it becomes a padding lane, a copy of A with weight zero (no branch around
        // the arithmetic: the 70 accumulator registers would have to be merged behind it)
        const int idxA = __float_as_int(mA.x);
        int idxB = __float_as_int(mB.x);
        float sqB = mB.y;
        if (idxB < 0) {
          xiB = xiA;
          xjB = xjA;
          idxB = idxA;
          sqB = 0.0f;
        }
        pair_accumulate<KIND, ROBUST>(acc2, c, pp, xiA, xiB, xjA, xjB, idxA, idxB, pk(mA.y, sqB));
        ra = rb;
        rb = rc;
        slot ^= 1;
      }
      cp_async_wait<0>();  // drain before the next segment reuses the ring
    }
    float acc[kAcc];
#pragma unroll
    for (int i = 0; i < kAcc; i++) acc[i] = lo(acc2[i]) + hi(acc2[i]);
    // ---- block reduction: transposing butterfly for acc[0..31] (lane l ends up with the warp total of acc[l]),
    //      plain xor-tree for the remaining kAcc-32 values, then the warps through shared memory
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
      partial[(size_t)(e + (int)blockIdx.x) * kAccStride + threadIdx.x] = v;
    }
    p = seg_end;
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
  const long long total = (long long)ws.L.num_kf * ws.L.num_pts;
  if (total == 0) return M3R_OK;
  pack_points_kernel<<<(unsigned)ceil_div_ll(total, 256), 256, 0, st>>>(Xs, Cs, ws.xc(), total, C_thresh,
                                                                        ws.state() + kCFail);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int launch_pack_edges(const GnWs& ws, const m3r_gn_params& p, const float* Cs, const int64_t* idx, const uint8_t* valid,
                      const float* Q, cudaStream_t st) {
  const int E = ws.L.num_edges, tiles = ws.L.tiles;
  M3R_CUDA(cudaMemsetAsync(ws.ecount(), 0, sizeof(int) * (size_t)(E + 1), st));
  if (E > 0 && tiles > 0) {
    M3R_CUDA(cudaMemsetAsync(ws.tstate(), 0, sizeof(int) * ((size_t)E * tiles + 2), st));
    pack_edges_kernel<<<(unsigned)((long long)E * tiles), 256, 0, st>>>(Cs, idx, valid, Q, ws.ie(), ws.je(), ws.rec(),
                                                                          ws.sqv(), ws.ecount(), ws.tstate(),
                                                                          ws.state() + kCFail, ws.L.num_pts, tiles,
                                                                          p.C_thresh, p.Q_thresh);
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
        Twc, ws.Kdev(), ws.xc(), ws.rec(), ws.sqv(), ws.eoff(), ws.ie(), ws.je(), ws.partial(), ws.state(), ws.L.num_pts,
        ws.L.num_edges, pp);
  else
    hessian_kernel<KIND, M3R_ROBUST_HUBER><<<grid, kHessThreads, sizeof(HessRing), st>>>(
        Twc, ws.Kdev(), ws.xc(), ws.rec(), ws.sqv(), ws.eoff(), ws.ie(), ws.je(), ws.partial(), ws.state(), ws.L.num_pts,
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
