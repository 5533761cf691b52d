// On-device linear algebra of the Gauss-Newton loop for sm_100a: keyframe index prep, dense fp64 assembly of the
// block system, blocked right-looking Cholesky with the right-hand side carried as an extra row, back substitution,
// Sim(3) retraction and the convergence flag.  Replaces the reference's host path: torch::_unique/searchsorted
// (gn_kernels.cu:161-170), SparseBlock + Eigen::SimplicialLLT in fp64 on the CPU with a D2H/H2D round trip and a
// .item() sync per iteration (:57-159, :1199-1222), and pose_retr_kernel (:415-453).
#include "gn_sim3.cuh"
#include "nd_plan.cuh"

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
// entry = 2*e + role (role 0: pose is the i end, 1: the j end).  inc_other[l] = free index o of the other end of entry
// l, with bit 30 set when l is the first entry of its row a that points at o and o < a (the entry that owns the
// off-diagonal block (a, o) in the assembly); -1 when the other end is the fixed pose.  Single CTA.
__global__ void __launch_bounds__(1024) incidence_kernel(const int* __restrict__ ie, const int* __restrict__ je,
                                                         int num_edges, int n_free, int* __restrict__ inc_off,
                                                         int* __restrict__ inc_list, int* __restrict__ inc_other) {
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
    // other ends + ownership flags of this row (this thread wrote the whole row, so no barrier is needed)
    const int beg = s_deg[a];
    for (int l = beg; l < o; l++) {
      const int ent = inc_list[l];
      const int e = ent >> 1;
      int other = ((ent & 1) ? ie[e] : je[e]) - 1;
      if (other >= 0 && other < a) {
        bool first = true;
        for (int l2 = beg; l2 < l; l2++) first = first && ((inc_other[l2] & 0x3fffffff) != other || inc_other[l2] < 0);
        if (first) other |= 1 << 30;
      }
      inc_other[l] = other;
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

__global__ void pose_retr_kernel(float* __restrict__ poses, const float* __restrict__ dx, int num_poses, int num_fix) {
  const int k = num_fix + blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= num_poses) return;
  float xi[7], p[8];
  for (int i = 0; i < 7; i++) xi[i] = dx[(size_t)(k - num_fix) * 7 + i];
  for (int i = 0; i < 8; i++) p[i] = poses[(size_t)k * 8 + i];
  retr_sim3(xi, p);
  for (int i = 0; i < 8; i++) poses[(size_t)k * 8 + i] = p[i];
}

// ============================================================================================== launchers
// ---- all of the above in ONE single-CTA kernel for graphs whose id list fits in shared memory (the usual case):
//      bitonic sort of the concatenated ids, unique, rank = lower_bound (== torch::_unique + searchsorted,
//      gn_kernels.cu:161-170), incidence lists, state reset.  The separate kernels cost 150 us per solve at 512 edges
//      (five launches, O(count^2) scans through global memory); this one takes ~15 us.
constexpr int kPrepThreads = 1024;
constexpr int kPrepMaxCount = 16384;

__global__ void __launch_bounds__(kPrepThreads) index_prep_kernel(const int64_t* __restrict__ all_ii,
                                                                  const int64_t* __restrict__ all_jj, int num_all,
                                                                  const int64_t* __restrict__ ii,
                                                                  const int64_t* __restrict__ jj, int num_edges,
                                                                  int num_kf, int padded, int* __restrict__ ie,
                                                                  int* __restrict__ je, int* __restrict__ inc_off,
                                                                  int* __restrict__ inc_list,
                                                                  int* __restrict__ inc_other, int* __restrict__ state,
                                                                  int* __restrict__ sky_first) {
  extern __shared__ __align__(16) unsigned char s_prep[];
  long long* ids = reinterpret_cast<long long*>(s_prep);                 // [padded]
  int* s_ie = reinterpret_cast<int*>(ids + padded);                       // [num_edges]
  int* s_je = s_ie + num_edges;                                           // [num_edges]
  int* s_deg = s_je + num_edges;                                          // [num_kf + 1]
  int* s_first = s_deg + num_kf + 1;                                      // [num_kf + 1]
  int* s_adj = s_first + num_kf + 1;                                      // [4 * nblk] tile adjacency bits (SkyPlan::adj)
  __shared__ int s_warp[kPrepThreads / 32];
  __shared__ int s_nuniq, s_bad;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long kInf = 0x7fffffffffffffffLL;
  const int count = 2 * num_all;
  if (tid < kStateInts) state[tid] = 0;
  if (tid == 0) s_bad = 0;
  for (int k = tid; k < padded; k += kPrepThreads)
    ids[k] = k < num_all ? (long long)all_ii[k] : (k < count ? (long long)all_jj[k - num_all] : kInf);
  __syncthreads();
  for (int size = 2; size <= padded; size <<= 1)
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = tid; t < (padded >> 1); t += kPrepThreads) {
        const int lo = 2 * t - (t & (stride - 1));
        const int hi = lo + stride;
        const bool up = (lo & size) == 0;
        const long long a = ids[lo], b = ids[hi];
        if ((a > b) == up) {
          ids[lo] = b;
          ids[hi] = a;
        }
      }
      __syncthreads();
    }
  // unique, compacted in place: every thread reads its chunk first, the writes (always to the left) come after a barrier
  const int per = (padded + kPrepThreads - 1) / kPrepThreads;  // <= 16
  long long mine[16];
  int nmine = 0;
  for (int q = 0; q < per; q++) {
    const int i = tid * per + q;
    if (i < padded) {
      const long long v = ids[i];
      if (v != kInf && (i == 0 || ids[i - 1] != v)) mine[nmine++] = v;
    }
  }
  int incl = nmine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = s_warp[lane];
    int wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += v;
    }
    s_warp[lane] = wi - w;
    if (lane == 31) s_nuniq = wi;
  }
  __syncthreads();
  {
    int pos = s_warp[warp] + incl - nmine;
    for (int q = 0; q < nmine; q++) ids[pos++] = mine[q];
  }
  __syncthreads();
  const int nuniq = s_nuniq;
  // rank of every local edge end
  for (int t = tid; t < 2 * num_edges; t += kPrepThreads) {
    const int e = t >> 1;
    const long long v = (t & 1) ? (long long)jj[e] : (long long)ii[e];
    int lo = 0, hi = nuniq;  // lower_bound
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (ids[mid] < v) lo = mid + 1; else hi = mid;
    }
    int r = lo;
    if (r >= num_kf) {  // more distinct keyframe ids than rows of Twc/Xs: the reference would read out of bounds
      r = -1;
      atomicAdd(&s_bad, 1);
    }
    if (t & 1) {
      je[e] = r;
      s_je[e] = r;
    } else {
      ie[e] = r;
      s_ie[e] = r;
    }
  }
  // envelope of the assembled system over ALL ranks' edges (the all-reduced [H | b] has the pattern of the whole graph):
  // first[a] = smallest free pose adjacent to free pose a, or a itself; consumed by skyline_plan_kernel
  const int n_free_env = num_kf > 0 ? num_kf - 1 : 0;
  const int nblk_env = (kPoseDim * n_free_env + kBlk - 1) / kBlk;
  const bool adj_on = nblk_env <= kSkyMaxBlk;
  for (int a = tid; a < n_free_env; a += kPrepThreads) s_first[a] = a;
  if (adj_on)
    for (int i = tid; i < 4 * nblk_env; i += kPrepThreads) s_adj[i] = 0;
  __syncthreads();
  auto mark_tile = [&](int ra, int rc) {
    const int hi = max(ra, rc), lo = min(ra, rc);
    atomicOr(&s_adj[4 * hi + (lo >> 5)], 1 << (lo & 31));
  };
  if (adj_on)
    for (int a = tid; a < n_free_env; a += kPrepThreads) {  // a pose's own 7 x 7 block (may straddle two row blocks)
      const int r0 = (kPoseDim * a) / kBlk, r1 = (kPoseDim * a + kPoseDim - 1) / kBlk;
      mark_tile(r0, r0);
      mark_tile(r1, r0);
      mark_tile(r1, r1);
    }
  for (int k = tid; k < num_all; k += kPrepThreads) {
    int r[2];
#pragma unroll
    for (int q = 0; q < 2; q++) {
      const long long v = q ? (long long)all_jj[k] : (long long)all_ii[k];
      int lo = 0, hi = nuniq;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (ids[mid] < v) lo = mid + 1; else hi = mid;
      }
      r[q] = lo;
    }
    if (r[0] >= num_kf || r[1] >= num_kf) continue;
    const int a = max(r[0], r[1]) - 1, o = min(r[0], r[1]) - 1;
    if (o >= 0) {
      atomicMin(&s_first[a], o);
      if (adj_on) {
        const int a0 = (kPoseDim * a) / kBlk, a1 = (kPoseDim * a + kPoseDim - 1) / kBlk;
        const int o0 = (kPoseDim * o) / kBlk, o1 = (kPoseDim * o + kPoseDim - 1) / kBlk;
        mark_tile(a0, o0);
        mark_tile(a0, o1);
        mark_tile(a1, o0);
        mark_tile(a1, o1);
      }
    }
  }
  __syncthreads();
  for (int a = tid; a < n_free_env; a += kPrepThreads) sky_first[a] = s_first[a];
  if (adj_on)
    for (int i = tid; i < 4 * nblk_env; i += kPrepThreads) sky_first[SkyPlan::adj(n_free_env) + i] = s_adj[i];
  if (tid == 0) state[kBadEdges] = s_bad;
  // incidence lists (see incidence_kernel): degrees with shared-memory atomics over all edge ends, offsets by one
  // thread, then one WARP per free keyframe walks the edges 32 at a time and places its entries with ballots, in the
  // order (edge, i-end before j-end) of the serial version (which spent ~40 us here with 31 active threads; r1)
  const int n_free = num_kf > 0 ? num_kf - 1 : 0;
  for (int a = tid; a <= n_free; a += kPrepThreads) s_deg[a] = 0;
  __syncthreads();
  for (int t = tid; t < 2 * num_edges; t += kPrepThreads) {
    const int e = t >> 1;
    const int i = s_ie[e], j = s_je[e];
    if (i < 0 || j < 0) continue;
    const int a = ((t & 1) ? j : i) - 1;
    if (a >= 0) atomicAdd(&s_deg[a], 1);
  }
  __syncthreads();
  if (tid == 0) {
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
  const unsigned lt = (1u << lane) - 1u;
  for (int a = warp; a < n_free; a += kPrepThreads / 32) {
    const int beg = s_deg[a];
    int o = beg;
    for (int e0 = 0; e0 < num_edges; e0 += 32) {
      const int e = e0 + lane;
      int i = -1, j = -1;
      if (e < num_edges) {
        i = s_ie[e];
        j = s_je[e];
      }
      const bool ok = i >= 0 && j >= 0;
      const bool mi = ok && (i - 1 == a), mj = ok && (j - 1 == a);
      const unsigned bi = __ballot_sync(0xffffffffu, mi), bj = __ballot_sync(0xffffffffu, mj);
      int pos = o + __popc(bi & lt) + __popc(bj & lt);
      if (mi) inc_list[pos++] = 2 * e;
      if (mj) inc_list[pos] = 2 * e + 1;
      o += __popc(bi) + __popc(bj);
    }
    __syncwarp();
    // the other end of every entry; bit 30 marks the first entry of each distinct lower neighbour (the owner of the
    // off-diagonal block in the assembly)
    for (int l0 = beg; l0 < o; l0 += 32) {
      const int l = l0 + lane;
      if (l < o) {
        const int ent = inc_list[l];
        const int e = ent >> 1;
        inc_other[l] = ((ent & 1) ? s_ie[e] : s_je[e]) - 1;
      }
    }
    __syncwarp();
    int flag[4] = {0, 0, 0, 0};  // up to 128 entries per keyframe in registers; longer lists fall back to a loop below
    for (int l0 = beg, q = 0; l0 < o; l0 += 32, q++) {
      const int l = l0 + lane;
      bool first = false;
      if (l < o) {
        const int other = inc_other[l];
        if (other >= 0 && other < a) {
          first = true;
          for (int l2 = beg; l2 < l; l2++) first = first && (inc_other[l2] != other);
        }
      }
      if (q < 4) flag[q] = first;  // entries beyond 128: see the serial tail
    }
    __syncwarp();
    for (int l0 = beg, q = 0; l0 < o && q < 4; l0 += 32, q++) {
      const int l = l0 + lane;
      if (l < o && (flag[q] & 1)) inc_other[l] |= 1 << 30;
    }
    if (o - beg > 128) {  // very high degree: finish serially (reads only raw values beyond entry 128)
      __syncwarp();
      if (lane == 0)
        for (int l = beg + 128; l < o; l++) {
          const int other = inc_other[l];
          if (other >= 0 && other < a) {
            bool first = true;
            for (int l2 = beg; l2 < l; l2++) first = first && ((inc_other[l2] & 0x3fffffff) != other || inc_other[l2] < 0);
            if (first) inc_other[l] = other | (1 << 30);
          }
        }
    }
  }
}

// ---- envelope plan (see SkyPlan, gn_internal.cuh).  One CTA, once per solve.  state[kSkyLanes] = lanes needed, or 0
//      when more than kSkyMaxLanes row blocks would be active at some step (the dense multi-CTA solver runs then).
int g_solver_mode = 0;  // 0: envelope solver where it applies, 1: always the dense multi-CTA solver (m3r_debug_set_solver)

__global__ void __launch_bounds__(128) skyline_plan_kernel(int* __restrict__ sky, int n_free, int nblk,
                                                           int* __restrict__ state) {
  __shared__ int s_fb[kSkyMaxBlk], s_lane[kSkyMaxBlk], s_over;
  const int tid = threadIdx.x;
  const int* first = sky + SkyPlan::first(n_free);
  if (tid == 0) s_over = 0;
  for (int rb = tid; rb < nblk; rb += blockDim.x) s_fb[rb] = rb;
  __syncthreads();
  for (int a = tid; a < n_free; a += blockDim.x) {
    const int cb = (kPoseDim * first[a]) / kBlk;  // first column block of the 7 rows of pose a
    const int r0 = (kPoseDim * a) / kBlk, r1 = (kPoseDim * a + kPoseDim - 1) / kBlk;
    atomicMin(&s_fb[r0], cb);
    if (r1 != r0) atomicMin(&s_fb[r1], cb);
  }
  __syncthreads();
  // schedule of step kb: the active row blocks in ascending order, members that join at this step flagged
  int* sched = sky + SkyPlan::sched(n_free);
  for (int kb = tid; kb < nblk; kb += blockDim.x) {
    int m = 0, mask = 0;
    for (int rb = kb; rb < nblk; rb++) {
      if (s_fb[rb] > kb) continue;
      if (m < kSkyMaxLanes) {
        sched[kb * kSkySched + 2 + m] = rb;
        if (s_fb[rb] == kb || (rb == kb && kb == 0)) mask |= 1 << m;
      }
      m++;
    }
    if (m > kSkyMaxLanes) atomicMax(&s_over, m);
    sched[kb * kSkySched] = m;
    sched[kb * kSkySched + 1] = mask;
  }
  __syncthreads();
  if (tid == 0) {
    // lanes: a row block takes the free lane that has been idle longest when it joins (step fb[rb]) and gives it back
    // after its own elimination.  If every lane taken at step kb was already idle during step kb - 1, the kernel may
    // fetch the new tiles one step early (bit 8 of the step's mask word).
    int used = 0, lanes = 0;
    int freed_at[kSkyMaxLanes];
    for (int l = 0; l < kSkyMaxLanes; l++) freed_at[l] = -4 - l;  // ties: lowest lane first
    if (!s_over) {
      for (int kb = 0; kb < nblk; kb++) {
        bool early = kb > 0;
        for (int rb = kb; rb < nblk; rb++) {
          if (s_fb[rb] != kb) continue;
          int best = -1;
          for (int l = 0; l < kSkyMaxLanes; l++)
            if (!(used & (1 << l)) && (best < 0 || freed_at[l] < freed_at[best])) best = l;
          used |= 1 << best;
          s_lane[rb] = best;
          if (best + 1 > lanes) lanes = best + 1;
          early = early && freed_at[best] <= kb - 2;
        }
        if (early && (sched[kb * kSkySched + 1] & 0xff)) sched[kb * kSkySched + 1] |= 1 << 8;
        used &= ~(1 << s_lane[kb]);
        freed_at[s_lane[kb]] = kb;
      }
    }
    state[kSkyLanes] = s_over ? 0 : lanes;
  }
  __syncthreads();
  for (int rb = tid; rb < nblk; rb += blockDim.x) {
    sky[SkyPlan::fb(n_free) + rb] = s_fb[rb];
    sky[SkyPlan::lane(n_free) + rb] = s_lane[rb];
  }
}

// ---- segmented schedule of the envelope solver (nd_plan.cuh), once per solve, after skyline_plan_kernel.  The symbolic
//      elimination is serial integer code (~0.18 ms at 256 keyframes): one candidate (interiors x separator width) per
//      CTA, one thread each, every valid candidate writes its whole plan into its own slot, the cheapest one is picked with
//      an atomic minimum and the CTA that finishes last publishes it if it beats the plain envelope solver (nd_accept):
//      state[kNdSegs] = number of interiors (0 = not segmented), state[kNdSlot] = the slot.  (As lanes of one warp the 21
//      candidates diverged and ran one after the other: 1.0 ms.)
__global__ void __launch_bounds__(32) nd_plan_kernel(int* __restrict__ sky, int n_free, int nblk, int* __restrict__ state,
                                                     int* __restrict__ ticket /* [0] best cost, [1] CTAs done */) {
  if (threadIdx.x != 0) return;
  if (state[kSkyLanes] > 0) {
    int P, sw, c16;
    nd_candidate(blockIdx.x, P, sw);
    int* plan = sky + SkyPlan::total(n_free) + (size_t)blockIdx.x * NdPlan::total;
    if (nd_simulate(sky + SkyPlan::adj(n_free), nblk, P, sw, plan, &c16)) atomicMin(ticket, c16 * 32 + (int)blockIdx.x);
  }
  __threadfence();
  if (atomicAdd(ticket + 1, 1) == (int)gridDim.x - 1) {  // last CTA: every plan and the minimum are visible
    __threadfence();
    const int cost = atomicMin(ticket, 0x7fffffff);
    int segs = 0, slot = 0;
    if (cost < 0x7f000000 && nd_accept(cost >> 5, nblk)) {  // (the ticket starts at 0x7f7f7f7f: no valid candidate)
      slot = cost & 31;
      segs = sky[SkyPlan::total(n_free) + (size_t)slot * NdPlan::total + NdPlan::hdr];
    }
    state[kNdSegs] = segs;
    state[kNdSlot] = slot;
  }
}

// one non-blocking side stream per device for the solver plans (created on first use, kept for the process)
static cudaStream_t plan_side_stream() {
  static cudaStream_t side[kMaxDevices] = {};
  static bool tried[kMaxDevices] = {};
  const int d = current_device_slot();
  if (!tried[d]) {
    tried[d] = true;
    if (cudaStreamCreateWithFlags(&side[d], cudaStreamNonBlocking) != cudaSuccess) {
      side[d] = nullptr;
      (void)cudaGetLastError();
    }
  }
  return side[d];
}

// the caller's stream waits for the plans that launch_index_prep put on the side stream (no-op if there are none)
int join_solver_plans(cudaEvent_t plan_join, cudaStream_t st) {
  if (plan_join == nullptr) return M3R_OK;
  M3R_CUDA(cudaStreamWaitEvent(st, plan_join, 0));
  M3R_CUDA(cudaEventDestroy(plan_join));
  return M3R_OK;
}

int launch_index_prep(const GnWs& ws, const int64_t* local_ii, const int64_t* local_jj, const int64_t* all_ii,
                      const int64_t* all_jj, int num_all_edges, cudaStream_t st, cudaEvent_t* plan_join) {
  *plan_join = nullptr;
  const int E = ws.L.num_edges;
  const int count = 2 * num_all_edges;
  if (count <= kPrepMaxCount) {
    int padded = 2;
    while (padded < count) padded <<= 1;
    const size_t smem = sizeof(long long) * (size_t)padded +
                        sizeof(int) * (size_t)(2 * E + 2 * ws.L.num_kf + 4 + 4 * (ws.L.nblk <= kSkyMaxBlk ? ws.L.nblk : 0));
    if (smem <= 200 * 1024) {
      static DynSmemOptIn attr;
      {
        const int rc = ensure_dyn_smem(index_prep_kernel, smem, attr);
        if (rc != M3R_OK) return rc;
      }
      index_prep_kernel<<<1, kPrepThreads, smem, st>>>(all_ii, all_jj, num_all_edges, local_ii, local_jj, E,
                                                         ws.L.num_kf, padded, ws.ie(), ws.je(), ws.inc_off(),
                                                         ws.inc_list(), ws.inc_other(), ws.state(), ws.sky());
      M3R_LAUNCH_CHECK();
      if (g_solver_mode != 1 && ws.L.n > kSmallMaxN && ws.L.nblk <= kSkyMaxBlk) {
        // The solver plans (0.07 + 0.21 ms at 256 keyframes, serial integer code on a handful of threads) are first
        // needed by the first solver launch: they run on a side stream next to the packing pre-pass (0.46 ms and more),
        // forked here and joined by join_solver_plans() at the end of gn_begin.
        cudaStream_t ps = st;
        *plan_join = nullptr;
        cudaStream_t side = plan_side_stream();
        cudaEvent_t fork = nullptr;
        if (side != nullptr && cudaEventCreateWithFlags(&fork, cudaEventDisableTiming) == cudaSuccess &&
            cudaEventCreateWithFlags(plan_join, cudaEventDisableTiming) == cudaSuccess) {
          M3R_CUDA(cudaEventRecord(fork, st));
          M3R_CUDA(cudaStreamWaitEvent(side, fork, 0));
          ps = side;
        } else {  // no side stream or no events: the plans simply run in line
          (void)cudaGetLastError();
          if (fork != nullptr) (void)cudaEventDestroy(fork);
          fork = nullptr;
          *plan_join = nullptr;
        }
        skyline_plan_kernel<<<1, 128, 0, ps>>>(ws.sky(), ws.L.n_free, ws.L.nblk, ws.state());
        M3R_LAUNCH_CHECK();
        if (g_solver_mode != 2 && ws.L.off_nd != 0 && ws.L.nblk >= 8) {
          int* ticket = ws.sky() + SkyPlan::total(ws.L.n_free) + kNdCandP * kNdCandSw * NdPlan::total;
          M3R_CUDA(cudaMemsetAsync(ticket, 0x7f, sizeof(int), ps));
          M3R_CUDA(cudaMemsetAsync(ticket + 1, 0, sizeof(int), ps));
          nd_plan_kernel<<<kNdCandP * kNdCandSw, 32, 0, ps>>>(ws.sky(), ws.L.n_free, ws.L.nblk, ws.state(), ticket);
          M3R_LAUNCH_CHECK();
        }
        if (ps != st) M3R_CUDA(cudaEventRecord(*plan_join, side));
        if (fork != nullptr) M3R_CUDA(cudaEventDestroy(fork));  // released once the work that waits on it has passed
      }
      return M3R_OK;
    }
  }
  init_state_kernel<<<1, 32, 0, st>>>(ws.state());
  M3R_LAUNCH_CHECK();
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
    incidence_kernel<<<1, 1024, smem, st>>>(ws.ie(), ws.je(), E, ws.L.n_free, ws.inc_off(), ws.inc_list(),
                                            ws.inc_other());
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

int launch_pose_retr(float* poses, const float* dx, int num_poses, int num_fix, cudaStream_t st) {
  const int cnt = num_poses - num_fix;
  if (cnt <= 0) return M3R_OK;
  pose_retr_kernel<<<ceil_div(cnt, 128), 128, 0, st>>>(poses, dx, num_poses, num_fix);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

}  // namespace m3r

// Host-side run of the planner (no GPU needed): the CPU tests check the schedule on block structures of their own.
// adj: [nblk][4] ints, bit cb of row rb (cb <= rb) = tile (rb, cb) structurally non-zero.  P = 0: pick like the device
// planner; otherwise simulate candidate (P, sw).  plan_out: NdPlan::total ints (m3r_debug_nd_plan_ints()).  Returns the
// number of interiors of the plan written, 0 if the structure does not split (or the candidate is invalid).
extern "C" int m3r_debug_nd_plan_ints(void) { return m3r::NdPlan::total; }
extern "C" int m3r_debug_nd_plan(const int* adj, int nblk, int P, int sw, int* plan_out) {
  using namespace m3r;
  if (adj == nullptr || plan_out == nullptr || nblk < 1 || nblk > kSkyMaxBlk) return 0;
  int c16 = 0;
  if (P > 0) return nd_simulate(adj, nblk, P, sw, plan_out, &c16) ? P : 0;
  int best = 0x7fffffff;
  for (int c = 0; c < kNdCandP * kNdCandSw; c++) {
    int p, w;
    nd_candidate(c, p, w);
    if (nd_simulate(adj, nblk, p, w, nullptr, &c16) && c16 * 32 + c < best) best = c16 * 32 + c;
  }
  if (best == 0x7fffffff || !nd_accept(best >> 5, nblk)) return 0;
  int p, w;
  nd_candidate(best & 31, p, w);
  return nd_simulate(adj, nblk, p, w, plan_out, &c16) ? p : 0;
}
