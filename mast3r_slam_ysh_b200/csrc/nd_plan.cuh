// Segmented (nested-dissection) schedule of the envelope Cholesky: which 32-row blocks are eliminated by which CTA.
//
// The single-CTA envelope solver (gn_fused.cu, sky_factor) walks the block columns of the assembled system in their
// natural order: a chain of nblk steps of ~14 us (56 steps, 0.75 ms, at 256 keyframes), replicated on every rank of a
// multi-GPU run.  A keyframe graph that is a chain or ring with a bounded window splits: pick P groups of `sw`
// consecutive row blocks as SEPARATORS, sw 32-row blocks being at least the block bandwidth, so that the interiors
// between them are not coupled directly.  The interiors are eliminated at the same time by P CTAs (each one a partial
// envelope factorisation that keeps its two neighbouring separators as long rows and accumulates their Schur complement
// in private tiles), one CTA sums the Schur complements in a fixed order and eliminates the separators, and the
// interiors back-substitute in parallel again.  The chain drops from nblk to nblk / P + P sw steps.  The matrix stays in
// its natural layout: a separator is a range of ROWS (it may cut through a keyframe's 7 unknowns), the permutation is only
// the order of the pivots, and a tile of L whose row block precedes its column block is kept in the (otherwise unused)
// upper triangle of sys.  The reference orders with AMD inside Eigen's SimplicialLLT (gn_kernels.cu:57-159) for the same
// reason: less fill, shorter elimination tree.
//
// Everything here is plain integer code shared by the device planner (gn_solve.cu, once per solve) and a host entry
// point (m3r_debug_nd_plan) that the CPU tests call.
#pragma once
#include "gn_internal.cuh"

namespace m3r {

constexpr int kNdMaxSeg = 8;     // interiors (CTAs); stage kNdMaxSeg at most is the separator stage
constexpr int kNdMaxBorder = 4;  // separator blocks alive at the end of an interior (two neighbours of two blocks)
constexpr int kNdBorderTiles = kNdMaxBorder * (kNdMaxBorder + 1) / 2;

// int offsets inside the plan region (workspace: sky + SkyPlan::total(n_free))
struct NdPlan {
  static constexpr int hdr = 0;                                        // [8] P (0: off), separator blocks, lanes, sw, cost x 16, [5] = planner scratch (best candidate)
  static constexpr int stage_begin = 8;                                // [kNdMaxSeg + 2] into order
  static constexpr int order = stage_begin + kNdMaxSeg + 2;            // [kSkyMaxBlk] pivot blocks in elimination order
  static constexpr int pos = order + kSkyMaxBlk;                       // [kSkyMaxBlk] inverse of order
  static constexpr int stage_of = pos + kSkyMaxBlk;                    // [kSkyMaxBlk] stage that eliminates the block
  static constexpr int sched = stage_of + kSkyMaxBlk;                  // [kSkyMaxBlk][kSkySched] by PIVOT block, see SkyPlan
  static constexpr int lane = sched + kSkyMaxBlk * kSkySched;          // [kNdMaxSeg + 1][kSkyMaxBlk] lane of a block in a stage
  static constexpr int adjS = lane + (kNdMaxSeg + 1) * kSkyMaxBlk;     // [kSkyMaxBlk][4] tiles of the separator stage (lower)
  static constexpr int lrow = adjS + 4 * kSkyMaxBlk;                   // [kSkyMaxBlk][4] column blocks of L in block row rb
  static constexpr int inS = lrow + 4 * kSkyMaxBlk;                    // [4] separator blocks
  static constexpr int exp = inS + 4;                                  // [kNdMaxSeg][1 + kNdMaxBorder] count, border blocks
  static constexpr int total = exp + kNdMaxSeg * (1 + kNdMaxBorder);
};

// scratch of the segmented solver (doubles, FusedArgs::ndbuf)
struct NdBuf {
  static constexpr size_t schur = 0;  // [kNdMaxSeg][kNdBorderTiles][32][32] Schur complement of interior s on its borders
  static constexpr size_t srhs = schur + (size_t)kNdMaxSeg * kNdBorderTiles * kBlk * kBlk;  // [kNdMaxSeg][kNdMaxBorder][32]
  static constexpr size_t x = srhs + (size_t)kNdMaxSeg * kNdMaxBorder * kBlk;                // [kSkyMaxBlk * 32] solution
  static constexpr size_t total = x + (size_t)kSkyMaxBlk * kBlk;
};

__host__ __device__ inline int nd_popc(unsigned x) {
#ifdef __CUDA_ARCH__
  return __popc(x);
#else
  return __builtin_popcount(x);
#endif
}
__host__ __device__ inline int nd_ffs(unsigned x) {  // index of the lowest set bit, x != 0
#ifdef __CUDA_ARCH__
  return __ffs((int)x) - 1;
#else
  return __builtin_ctz(x);
#endif
}
__host__ __device__ inline bool nd_bit(const unsigned* m, int i) { return (m[i >> 5] >> (i & 31)) & 1u; }
__host__ __device__ inline void nd_set(unsigned* m, int i) { m[i >> 5] |= 1u << (i & 31); }
__host__ __device__ inline void nd_clr(unsigned* m, int i) { m[i >> 5] &= ~(1u << (i & 31)); }

// Geometry of candidate (P, sw): interiors of near-equal length, each followed by sw separator blocks; the last separator
// is the tail of the matrix (for a ring it carries the wrap-around rows, for an open chain it is merely a small last stage).
__host__ __device__ inline void nd_geometry(int nblk, int P, int sw, signed char* stage_of) {
  const int nI = nblk - P * sw;
  int b = 0;
  for (int p = 0; p < P; p++) {
    const int q = nI / P + (p < nI % P ? 1 : 0);
    for (int i = 0; i < q; i++) stage_of[b++] = (signed char)p;
    for (int i = 0; i < sw; i++) stage_of[b++] = (signed char)P;
  }
}

// Symbolic elimination of candidate (P, sw) on the block structure `adj` ([nblk][4], bit cb of row rb, cb <= rb: tile
// (rb, cb) of the assembled matrix is structurally non-zero).  Returns false when the candidate is not a valid dissection
// (two interiors coupled, more than kSkyMaxLanes active row blocks, more than kNdMaxBorder separator blocks per
// interior).  *cost16 = 16 x the estimated chain length in steps.  With out != nullptr the plan is written there.
__host__ __device__ inline bool nd_simulate(const int* adj, int nblk, int P, int sw, int* out, int* cost16) {
  if (P < 2 || P > kNdMaxSeg || sw < 1 || nblk > kSkyMaxBlk || nblk - P * sw < P) return false;
  unsigned S[kSkyMaxBlk][4];  // symmetric structure, fill included as the elimination proceeds
  signed char stage_of[kSkyMaxBlk];
  nd_geometry(nblk, P, sw, stage_of);
  for (int rb = 0; rb < nblk; rb++)
    for (int w = 0; w < 4; w++) S[rb][w] = (unsigned)adj[rb * 4 + w];
  for (int rb = 0; rb < nblk; rb++) {
    nd_set(S[rb], rb);
    for (int w = 0; w < 4; w++) {
      unsigned m = (unsigned)adj[rb * 4 + w];
      while (m) {
        const int cb = 32 * w + nd_ffs(m);
        m &= m - 1;
        if (cb < rb) nd_set(S[cb], rb);
      }
    }
  }
  unsigned inS[4] = {0, 0, 0, 0}, E[4] = {0, 0, 0, 0};
  int nS = 0;
  for (int rb = 0; rb < nblk; rb++)
    if (stage_of[rb] == P) {
      nd_set(inS, rb);
      nS++;
    }
  if (out) {
    for (int i = 0; i < NdPlan::total; i++) out[i] = 0;
    for (int w = 0; w < 4; w++) out[NdPlan::inS + w] = (int)inS[w];
    for (int rb = 0; rb < nblk; rb++) out[NdPlan::stage_of + rb] = stage_of[rb];
  }
  int lanes = 0, longest = 0, npiv = 0;
  for (int s = 0; s <= P; s++) {
    unsigned A[4] = {0, 0, 0, 0};
    int used = 0, t = 0, len = 0;
    int freed_at[kSkyMaxLanes];
    signed char lane_of[kSkyMaxBlk];
    for (int l = 0; l < kSkyMaxLanes; l++) freed_at[l] = -4 - l;  // ties: lowest lane first
    if (out) out[NdPlan::stage_begin + s] = npiv;
    for (int k = 0; k < nblk; k++) {
      if (stage_of[k] != s) continue;
      // members of pivot k: its structure minus what has been eliminated
      unsigned N[4], nw[4];
      int cnt = 0;
      for (int w = 0; w < 4; w++) {
        N[w] = S[k][w] & ~E[w];
        nw[w] = N[w] & ~A[w];
        A[w] |= N[w];
        cnt += nd_popc(A[w]);
      }
      if (cnt > kSkyMaxLanes) return false;
      // a member must belong to this interior or be a separator block
      int mem[kSkyMaxLanes], m = 0;
      mem[m++] = k;
      for (int pass = 0; pass < 2; pass++)  // elimination order: the interior's own blocks, then the separators
        for (int w = 0; w < 4; w++) {
          unsigned a = A[w];
          while (a) {
            const int r = 32 * w + nd_ffs(a);
            a &= a - 1;
            if (r == k) continue;
            const int sr = stage_of[r];
            if (sr != s && sr != P) return false;
            if (s < P && sr == s && r < k) return false;  // cannot happen (eliminated blocks leave A); defensive
            if ((pass == 0) == (sr == s && s < P)) mem[m++] = r;
          }
        }
      if (s == P) {  // separator stage: plain ascending order (pass 1 collected everything; keep it sorted)
        for (int i = 1; i < m; i++)
          for (int j = i; j > 1 && mem[j] < mem[j - 1]; j--) {
            const int x = mem[j];
            mem[j] = mem[j - 1];
            mem[j - 1] = x;
          }
      }
      // lanes of the new members
      int mask = 0;
      bool early = t > 0;
      for (int i = 0; i < m; i++) {
        const int r = mem[i];
        if (!nd_bit(nw, r)) continue;
        int best = -1;
        for (int l = 0; l < kSkyMaxLanes; l++)
          if (!(used & (1 << l)) && (best < 0 || freed_at[l] < freed_at[best])) best = l;
        if (best < 0) return false;
        used |= 1 << best;
        lane_of[r] = (signed char)best;
        if (best + 1 > lanes) lanes = best + 1;
        early = early && freed_at[best] <= t - 2;
        mask |= 1 << i;
        if (out) out[NdPlan::lane + s * kSkyMaxBlk + r] = best;
      }
      if (early && mask) mask |= 1 << 8;
      if (out) {
        int* sc = out + NdPlan::sched + k * kSkySched;
        sc[0] = m;
        sc[1] = mask;
        for (int i = 0; i < m; i++) sc[2 + i] = mem[i];
        out[NdPlan::order + npiv] = k;
        out[NdPlan::pos + k] = npiv;
      }
      // fill among the structural members, rows of L
      for (int w = 0; w < 4; w++) {
        unsigned a = N[w];
        while (a) {
          const int r = 32 * w + nd_ffs(a);
          a &= a - 1;
          if (r == k) continue;
          for (int w2 = 0; w2 < 4; w2++) S[r][w2] |= N[w2];
          if (out) out[NdPlan::lrow + r * 4 + (k >> 5)] |= (int)(1u << (k & 31));
        }
      }
      nd_set(E, k);
      nd_clr(A, k);
      used &= ~(1 << lane_of[k]);
      freed_at[lane_of[k]] = t;
      t++;
      len++;
      npiv++;
    }
    if (s < P) {
      // what is left are separator blocks: this interior's Schur complement lives on them
      int nb = 0;
      for (int w = 0; w < 4; w++) {
        if (A[w] & ~inS[w]) return false;
        unsigned a = A[w];
        while (a) {
          const int r = 32 * w + nd_ffs(a);
          a &= a - 1;
          if (nb >= kNdMaxBorder) return false;
          if (out) out[NdPlan::exp + s * (1 + kNdMaxBorder) + 1 + nb] = r;
          nb++;
        }
      }
      if (out) out[NdPlan::exp + s * (1 + kNdMaxBorder)] = nb;
      if (len > longest) longest = len;
    }
  }
  if (out) {
    out[NdPlan::stage_begin + P + 1] = npiv;
    for (int rb = 0; rb < nblk; rb++) {
      if (!nd_bit(inS, rb)) continue;
      for (int w = 0; w < 4; w++) {
        unsigned low = S[rb][w] & inS[w];  // separator tiles at or left of the diagonal
        if (32 * w + 31 > rb) low &= (rb >= 32 * w) ? ((rb - 32 * w == 31) ? 0xffffffffu : ((1u << (rb - 32 * w + 1)) - 1u)) : 0u;
        out[NdPlan::adjS + rb * 4 + w] = (int)(low | ((unsigned)adj[rb * 4 + w] & inS[w]));
      }
    }
    out[NdPlan::hdr + 0] = P;
    out[NdPlan::hdr + 1] = nS;
    out[NdPlan::hdr + 2] = lanes;
    out[NdPlan::hdr + 3] = sw;
  }
  // Measured (tools/time_phases.py, rings and chains of 40-256 keyframes): a separator step costs about what an interior
  // step costs (fewer rows to solve, more tiles to update), and the three barriers, the reduce and the per-stage set-up
  // cost about three steps.  The plan is used if that beats the plain chain of nblk steps (nd_accept).
  *cost16 = 16 * longest + 16 * nS + 48;
  if (out) out[NdPlan::hdr + 4] = *cost16;
  return true;
}

__host__ __device__ inline bool nd_accept(int cost16, int nblk) { return cost16 < 16 * nblk; }

// candidates tried by the planner, the cheapest valid one wins if nd_accept says it beats the plain envelope solver
constexpr int kNdCandP = 7;   // P = 2 .. 8
constexpr int kNdCandSw = 3;  // sw = 1 .. 3
__host__ __device__ inline void nd_candidate(int c, int& P, int& sw) {
  P = 2 + c % kNdCandP;
  sw = 1 + c / kNdCandP;
}

}  // namespace m3r
