// Tile-staged refine_matches for sm_100a (opt-in variant of the per-thread gathers of matching.cu, for queries that form
// the h x w grid of an image, which is how every caller of the reference uses it: matching.py:79-85, n == h * w).
//
// refine_matches_kernel (matching_kernels.cu:25-81) scores 245 candidates of 48 bytes per query: 11.8 KB of scattered
// 16-byte loads per query through L1 (l1tex 97 % busy, profiles/r1_refine_v1_ncu_full.txt), although neighbouring queries
// look at nearly the same window of D11.  Here a CTA owns a 16 x 16 tile of query pixels.  Per dilation level it
//   1. takes the mean of its queries' current centres (one block reduction),
//   2. has ONE thread fetch the box of D11 around that mean with TMA tensor copies (cp.async.bulk.tensor, SASS UTMALDG;
//      48 x 48 pixels for the wide levels, 32 x 32 for the narrow ones; out-of-image pixels are zero-filled by the copy
//      engine), completion on an mbarrier.  The maps have the IMAGE ROW as innermost dimension and a box is fetched as
//      column groups of 16 pixels (16 x 48 B = 192 fp32 elements, the copy engine takes at most 256 per request): 144
//      requests per 48 x 48 box.  The first version described a box over {fdim, w, h, b}: one request per PIXEL, 2304 per
//      box and 11.5 k per CTA, and the copy engine's request rate (~10 cycles each), not the scoring, set the pace,
//   3. scores out of shared memory: every query whose whole (2 r d + 1)^2 window lies inside the box reads its candidates
//      with LDS.128 (48-byte pixel pitch: conflict-free for neighbouring pixels), the others -- stragglers whose centre
//      left the neighbourhood -- fall back to global loads,
//   4. recentres (matching_kernels.cu:75-76) and goes on to the next level.
// Arithmetic, scan order and the strict-greater argmax are those of matching_core.cuh (score_f16_multi, u offset outer, v
// offset inner), so the result is bit-identical to the gather kernel and to the reference.
#include <cuda_fp16.h>

#include "matching_core.cuh"
#include "tensormap.cuh"

namespace m3r {

constexpr int kRtQ = 16;        // query tile edge
constexpr int kRtThreads = kRtQ * kRtQ;
constexpr int kRtBoxBig = 48;   // box edge for levels with 16 + 2 r d <= 48
constexpr int kRtBoxSmall = 32;
constexpr int kRtGroup = 16;    // pixels per column group = one copy request per box row

template <int FDIM>
struct RefineTileSmem {
  uint4 tile[kRtBoxBig * kRtBoxBig * (FDIM / 8)];  // box as column groups of 16 pixels: [group][row][16 px][FDIM halves]
  unsigned long long bar;
  int red[3][kRtThreads / 32];
  int origin[3];  // ox, oy, box edge (0: no box this level)
};

template <int FDIM, int RADIUS>
__global__ void __launch_bounds__(kRtThreads, 2)
    refine_matches_tile_kernel(const __grid_constant__ CUtensorMap map_big, const __grid_constant__ CUtensorMap map_small,
                               const uint4* __restrict__ D11, const uint4* __restrict__ D21,
                               const longlong2* __restrict__ p1, longlong2* __restrict__ p1_new, int h, int w,
                               int dilation_max) {
  constexpr int NV = FDIM / 8;
  constexpr int NJ = 2 * RADIUS + 1;
  extern __shared__ __align__(128) unsigned char s_raw[];
  RefineTileSmem<FDIM>& sm =  // TMA destinations must be 128-byte aligned
      *reinterpret_cast<RefineTileSmem<FDIM>*>(s_raw + ((128u - (smem_u32(s_raw) & 127u)) & 127u));
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.z;
  const int qx = blockIdx.x * kRtQ + (tid & (kRtQ - 1)), qy = blockIdx.y * kRtQ + (tid >> 4);
  const bool has_query = qx < w && qy < h;
  const long long gid = (long long)b * h * w + (long long)qy * w + qx;
  if (tid == 0) mbar_init(&sm.bar, 1);

  longlong2 p = make_longlong2(0, 0);
  if (has_query) p = p1[gid];
  // coordinates far outside any image can never produce an in-image candidate: result == input
  const long long lim = 1LL << 29;
  const bool far_out = p.x < -lim || p.x > lim || p.y < -lim || p.y > lim;
  const bool active = has_query && !far_out;
  __half2 q[FDIM / 2];
  if (active) load_query_f16<FDIM>(D21, gid, q);
  const uint4* __restrict__ img = D11 + (size_t)b * NV * h * w;
  int u0 = active ? (int)p.x : 0, v0 = active ? (int)p.y : 0;
  int u_new = u0, v_new = v0;
  float max_score = 0.0f;  // cuda::std::numeric_limits<c10::Half>::min() == Half() == +0 (matching_kernels.cu:47)
  unsigned level = 0;
  __syncthreads();  // barrier initialised

  for (int d = dilation_max; d > 0; d--, level++) {
    const int rd = RADIUS * d;
    // ---- 1. mean centre of the tile's active queries
    {
      int su = active ? u0 : 0, sv = active ? v0 : 0, sc = active ? 1 : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        su += __shfl_xor_sync(0xffffffffu, su, o);
        sv += __shfl_xor_sync(0xffffffffu, sv, o);
        sc += __shfl_xor_sync(0xffffffffu, sc, o);
      }
      if (lane == 0) {
        sm.red[0][warp] = su;
        sm.red[1][warp] = sv;
        sm.red[2][warp] = sc;
      }
    }
    __syncthreads();  // also: every thread is done reading the previous level's box
    // ---- 2. one thread fetches the box
    if (tid == 0) {
      long long su = 0, sv = 0;
      int sc = 0;
#pragma unroll
      for (int k = 0; k < kRtThreads / 32; k++) {
        su += sm.red[0][k];
        sv += sm.red[1][k];
        sc += sm.red[2][k];
      }
      const int need = kRtQ + 2 * rd;
      const int box = need <= kRtBoxSmall ? kRtBoxSmall : kRtBoxBig;
      int ox = 0, oy = 0, edge = 0;
      if (sc > 0) {
        ox = (int)(su / sc) - box / 2;
        oy = (int)(sv / sc) - box / 2;
        edge = box;
        mbar_arrive_expect_tx(&sm.bar, (unsigned)(box * box * FDIM * 2));
        for (int g = 0; g < box / kRtGroup; g++)  // coordinates along the row in fp32 elements (FDIM / 2 per pixel)
          tma_load_4d(sm.tile + (size_t)g * box * kRtGroup * NV, box == kRtBoxSmall ? &map_small : &map_big,
                      (ox + kRtGroup * g) * (FDIM / 2), oy, b, 0, &sm.bar);
      }
      sm.origin[0] = ox;
      sm.origin[1] = oy;
      sm.origin[2] = edge;
    }
    __syncthreads();
    const int ox = sm.origin[0], oy = sm.origin[1], box = sm.origin[2];
    if (box == 0) break;  // no active query in this tile (uniform)
    const bool in_box = active && u0 - rd >= ox && u0 + rd < ox + box && v0 - rd >= oy && v0 + rd < oy + box;
    mbar_wait(&sm.bar, level & 1);
    // ---- 3. scan (matching_kernels.cu:51-69): u offset outer, v offset inner, the 7 candidates of a column together
    if (active) {
      for (int i = 0; i <= 2 * rd; i += d) {
        const int u = u0 - rd + i;
        if (u < 0 || u >= w) continue;
        uint4 t[NJ][NV];
        bool in[NJ];
        if (in_box) {
#pragma unroll
          for (int jj = 0; jj < NJ; jj++) {
            const int v = v0 - rd + jj * d;
            in[jj] = v >= 0 && v < h;
            const int lx = u - ox;
            const uint4* c = sm.tile + (size_t)(((lx >> 4) * box + (v - oy)) * kRtGroup + (lx & (kRtGroup - 1))) * NV;
#pragma unroll
            for (int k = 0; k < NV; k++) t[jj][k] = c[k];
          }
        } else {
#pragma unroll
          for (int jj = 0; jj < NJ; jj++) {
            const int v = v0 - rd + jj * d;
            in[jj] = v >= 0 && v < h;
            const uint4* c = img + (size_t)((in[jj] ? v : 0) * w + u) * NV;
#pragma unroll
            for (int k = 0; k < NV; k++) t[jj][k] = in[jj] ? __ldg(c + k) : make_uint4(0, 0, 0, 0);
          }
        }
        float sc[NJ];
        score_f16_multi<NV, NJ>(q, t, sc);  // the NJ rounding chains interleaved (matching_core.cuh)
#pragma unroll
        for (int jj = 0; jj < NJ; jj++) {
          if (in[jj] && sc[jj] > max_score) {
            max_score = sc[jj];
            u_new = u;
            v_new = v0 - rd + jj * d;
          }
        }
      }
      // ---- 4. recentre
      u0 = u_new;
      v0 = v_new;
    }
  }
  if (has_query) p1_new[gid] = far_out ? p : make_longlong2((long long)u_new, (long long)v_new);
}

template <int FDIM>
static int launch_tile_fdim(const CUtensorMap& big, const CUtensorMap& small, const uint16_t* D11, const uint16_t* D21,
                            const int64_t* p1, int64_t* p1_new, int b, int h, int w, int dilation_max, cudaStream_t st) {
  static DynSmemOptIn attr;  // per device
  const size_t smem = sizeof(RefineTileSmem<FDIM>) + 128;
  const int rc = ensure_dyn_smem(refine_matches_tile_kernel<FDIM, 3>, smem, attr);
  if (rc != M3R_OK) return rc;
  const dim3 grid((unsigned)ceil_div(w, kRtQ), (unsigned)ceil_div(h, kRtQ), (unsigned)b);
  refine_matches_tile_kernel<FDIM, 3><<<grid, kRtThreads, smem, st>>>(big, small, (const uint4*)D11, (const uint4*)D21,
                                                                        (const longlong2*)p1, (longlong2*)p1_new, h, w,
                                                                        dilation_max);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

// returns M3R_OK when the tile kernel ran, 1 when the caller should use the gather kernel, else an error code
int launch_refine_tile_f16(const uint16_t* D11, const uint16_t* D21, const int64_t* p1, int64_t* p1_new, int b, int h,
                           int w, int fdim, int radius, int dilation_max, cudaStream_t st) {
  if (radius != 3 || (fdim != 24 && fdim != 16) || dilation_max < 1) return 1;
  if (kRtQ + 2 * radius * dilation_max > kRtBoxBig) return 1;  // widest window must fit the big box
  if (b > 65535 || ((uintptr_t)D11 & 15) != 0) return 1;
  CUtensorMap big, small;
  // descriptor rows as fp32 elements (two halves each; the copy engine only moves bytes)
  if (!encode_row_tiles_f32(&big, D11, b, h, w * (fdim / 2), kRtGroup * (fdim / 2), kRtBoxBig)) return 1;
  if (!encode_row_tiles_f32(&small, D11, b, h, w * (fdim / 2), kRtGroup * (fdim / 2), kRtBoxSmall)) return 1;
  if (fdim == 24) return launch_tile_fdim<24>(big, small, D11, D21, p1, p1_new, b, h, w, dilation_max, st);
  return launch_tile_fdim<16>(big, small, D11, D21, p1, p1_new, b, h, w, dilation_max, st);
}

// ================================================================================================ iter_proj, tile-staged
// iter_proj_v2_kernel (matching.cu) is bound by L1 wavefronts: every LM iteration gathers four bilinear corners from
// three planes, and on real (noisy) pointmaps the 32 queries of a warp land on different rows of the ray image, so one
// LDG.128 touches up to 32 cache lines.  Queries that are neighbours in the image still converge to the same
// neighbourhood.  Here a CTA owns a 16 x 8 tile of queries (n == h * w, the callers' layout: matching.py:41-60) and
//   1. runs the first `warm` LM iterations from global memory (the first Gauss-Newton step is the long one: p_init may be
//      a hundred pixels from the solution),
//   2. takes the mean of its queries' positions and has ONE thread fetch the 32 x 24 pixel window around it from the
//      three repacked planes with TMA tensor copies (rank-4 maps with the image row innermost, so a window row is one
//      request; the window starts at a multiple of four pixels; 27 KB, one mbarrier),
//   3. runs the remaining iterations out of shared memory: a cell whose four corners lie in the window is read with
//      LDS.128 (at most 4 wavefronts per warp instead of one per row touched), any other cell falls back to global loads.
// The arithmetic is iter_proj_steps of matching_core.cuh in both phases: bit-identical to the gather kernel.
constexpr int kIpQW = 16, kIpQH = 8;
constexpr int kIpThreads = kIpQW * kIpQH;
constexpr int kIpBW = 32, kIpBH = 24;

struct IterProjTileSmem {
  float4 t0[kIpBH * kIpBW];
  float4 t1[kIpBH * kIpBW];
  float t2[kIpBH * kIpBW];
  unsigned long long bar;
  float red[3][kIpThreads / 32];
  int origin[2];
};

struct TileRaySampler {
  GlobalRaySampler g;
  const float4* t0;
  const float4* t1;
  const float* t2;
  int ox, oy;
#ifdef M3R_IP_STATS
  mutable unsigned hits = 0, misses = 0;
#endif
  __device__ __forceinline__ bool inside(const Bil& B, int& i) const {
    const unsigned lx = (unsigned)(B.u11 - ox), ly = (unsigned)(B.v11 - oy);
    i = (int)(ly * kIpBW + lx);
    const bool in = lx < (unsigned)(kIpBW - 1) && ly < (unsigned)(kIpBH - 1);
#ifdef M3R_IP_STATS
    if (in) hits++; else misses++;
#endif
    return in;
  }
  __device__ __forceinline__ void ray(const Bil& B, float4& a11, float4& a12, float4& a21, float4& a22) const {
    int i;
    if (inside(B, i)) {
      a22 = t0[i];
      a21 = t0[i + 1];
      a12 = t0[i + kIpBW];
      a11 = t0[i + kIpBW + 1];
    } else {
      g.ray(B, a11, a12, a21, a22);
    }
  }
  __device__ __forceinline__ void grad(const Bil& B, float4& c11, float4& c12, float4& c21, float4& c22, float& d11,
                                       float& d12, float& d21, float& d22) const {
    int i;
    if (inside(B, i)) {
      c22 = t1[i];
      c21 = t1[i + 1];
      c12 = t1[i + kIpBW];
      c11 = t1[i + kIpBW + 1];
      d22 = t2[i];
      d21 = t2[i + 1];
      d12 = t2[i + kIpBW];
      d11 = t2[i + kIpBW + 1];
    } else {
      g.grad(B, c11, c12, c21, c22, d11, d12, d21, d22);
    }
  }
};

__global__ void __launch_bounds__(kIpThreads)
    iter_proj_tile_kernel(const __grid_constant__ CUtensorMap map0, const __grid_constant__ CUtensorMap map1,
                          const __grid_constant__ CUtensorMap map2, const float4* __restrict__ P0,
                          const float4* __restrict__ P1, const float* __restrict__ P2, const float* __restrict__ pts,
                          const float* __restrict__ p_init, float* __restrict__ p_new, uint8_t* __restrict__ converged,
                          int h, int w, int max_iter, int warm, float lambda_init, float cost_thresh,
                          unsigned long long* __restrict__ stats) {
  extern __shared__ __align__(128) unsigned char s_raw[];
  IterProjTileSmem& sm =  // TMA destinations must be 128-byte aligned
      *reinterpret_cast<IterProjTileSmem*>(s_raw + ((128u - (smem_u32(s_raw) & 127u)) & 127u));
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.z;
  const int qx = blockIdx.x * kIpQW + (tid & (kIpQW - 1)), qy = blockIdx.y * kIpQH + (tid / kIpQW);
  const bool has = qx < w && qy < h;
  const size_t img = (size_t)b * h * w;
  const size_t gid = img + (size_t)qy * w + qx;
  if (tid == 0) mbar_init(&sm.bar, 1);

  float tx = 0.f, ty = 0.f, tz = 1.f;
  float2 pi = make_float2(1.f, 1.f);
  if (has) {
    pi = __ldg(reinterpret_cast<const float2*>(p_init) + gid);
    tx = __ldg(pts + gid * 3 + 0);
    ty = __ldg(pts + gid * 3 + 1);
    tz = __ldg(pts + gid * 3 + 2);
  }
  IterProjState s;
  iter_proj_begin(s, fminf(fmaxf(pi.x, 1.0f), (float)(w - 2)), fminf(fmaxf(pi.y, 1.0f), (float)(h - 2)), w, lambda_init);
  const GlobalRaySampler G{P0 + img, P1 + img, P2 + img, w};
  const int it_warm = min(warm, max_iter);
  if (has) iter_proj_steps(s, G, tx, ty, tz, w, h, 0, it_warm, cost_thresh);

  if (it_warm < max_iter) {  // uniform over the CTA
    // ---- mean position of the tile's queries
    float su = has ? s.u : 0.f, sv = has ? s.v : 0.f, sc = has ? 1.f : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      su += __shfl_xor_sync(0xffffffffu, su, o);
      sv += __shfl_xor_sync(0xffffffffu, sv, o);
      sc += __shfl_xor_sync(0xffffffffu, sc, o);
    }
    if (lane == 0) {
      sm.red[0][warp] = su;
      sm.red[1][warp] = sv;
      sm.red[2][warp] = sc;
    }
    __syncthreads();  // also: the mbarrier is initialised
    if (tid == 0) {
      float mu = 0.f, mv = 0.f, mc = 0.f;
#pragma unroll
      for (int k = 0; k < kIpThreads / 32; k++) {
        mu += sm.red[0][k];
        mv += sm.red[1][k];
        mc += sm.red[2][k];
      }
      // mc >= 1: a CTA always has its first query
      // the window starts at a multiple of four pixels (floor, also for negative origins): 16-byte aligned rows in all
      // three planes
      const int ox = (((int)(mu / mc) - (kIpBW / 2 - 1)) >> 2) << 2, oy = (int)(mv / mc) - (kIpBH / 2 - 1);
      sm.origin[0] = ox;
      sm.origin[1] = oy;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_arrive_expect_tx(&sm.bar, (unsigned)(sizeof(sm.t0) + sizeof(sm.t1) + sizeof(sm.t2)));
      tma_load_4d(sm.t0, &map0, 4 * ox, oy, b, 0, &sm.bar);  // coordinates in floats along the row
      tma_load_4d(sm.t1, &map1, 4 * ox, oy, b, 0, &sm.bar);
      tma_load_4d(sm.t2, &map2, ox, oy, b, 0, &sm.bar);
    }
    __syncthreads();
    TileRaySampler T{G, sm.t0, sm.t1, sm.t2, sm.origin[0], sm.origin[1]};
    mbar_wait(&sm.bar, 0);
    if (has) iter_proj_steps(s, T, tx, ty, tz, w, h, it_warm, max_iter, cost_thresh);
#ifdef M3R_IP_STATS
    if (stats != nullptr) {
      atomicAdd(stats + 0, (unsigned long long)T.hits);
      atomicAdd(stats + 1, (unsigned long long)T.misses);
    }
#endif
  }
  if (has) {
    reinterpret_cast<float2*>(p_new)[gid] = make_float2(s.u, s.v);
    converged[gid] = s.conv ? 1 : 0;
  }
}

static unsigned long long* g_ip_stats = nullptr;

// returns M3R_OK when the tile kernel ran, 1 when the caller should use the gather kernel, else an error code
int launch_iter_proj_tile(const float4* P0, const float4* P1, const float* P2, const float* pts, const float* p_init,
                          float* p_new, uint8_t* converged, int b, int h, int w, int max_iter, int warm,
                          float lambda_init, float cost_thresh, cudaStream_t st) {
  if (b > 65535 || b < 1 || (w & 3) != 0) return 1;
  CUtensorMap m0, m1, m2;
  // image rows are the innermost dimension of the maps: a window row is ONE request of the copy engine (512 / 512 / 128
  // bytes), 72 requests per window.  (First version: boxes over {4 floats, w, h, b}, a request per pixel, 2304 per
  // window -- the engine's request rate made the kernel slower than the gathers even with every corner served from the
  // window.  A rank-3 map over the scalar plane raised an illegal instruction on the box, hence rank 4 with a unit
  // dimension.)
  if (!encode_row_tiles_f32(&m0, P0, b, h, 4 * w, 4 * kIpBW, kIpBH)) return 1;
  if (!encode_row_tiles_f32(&m1, P1, b, h, 4 * w, 4 * kIpBW, kIpBH)) return 1;
  if (!encode_row_tiles_f32(&m2, P2, b, h, w, kIpBW, kIpBH)) return 1;
  const size_t smem = sizeof(IterProjTileSmem) + 128;
  const dim3 grid((unsigned)ceil_div(w, kIpQW), (unsigned)ceil_div(h, kIpQH), (unsigned)b);
  iter_proj_tile_kernel<<<grid, kIpThreads, smem, st>>>(m0, m1, m2, P0, P1, P2, pts, p_init, p_new, converged, h, w,
                                                        max_iter, warm, lambda_init, cost_thresh, g_ip_stats);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

}  // namespace m3r

// diagnostics (-DM3R_IP_STATS builds only): two device counters, corner fetches served by the window / by the fallback
extern "C" void m3r_debug_set_iter_proj_stats(unsigned long long* dev_counters) { m3r::g_ip_stats = dev_counters; }
