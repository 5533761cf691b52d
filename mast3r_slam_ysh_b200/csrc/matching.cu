// iter_proj and refine_matches for sm_100a.
//
// Replaces the reference's matching_kernels.cu (iter_proj_kernel :119-275, refine_matches_kernel :25-81) behind
// the C ABI of include/m3r_backends.h.  Written from the behaviour documented in SURVEY.md section 8(a):
//   * one thread per query point, flat 1-D grid over (batch x points) with a bounds guard (the reference uses
//     16-thread blocks without a guard),
//   * read-only (LDG.NC) loads so the ray image / descriptor map stay L1-resident across neighbouring queries,
//   * `converged` kept in a register and stored once (the reference stores it every LM iteration),
//   * fp16 scores use the non-contracting __hmul2_rn / __hadd_rn intrinsics: each product and each partial sum is
//     rounded to fp16, which is bit-identical to the reference's c10::Half arithmetic (fp32 op + round to fp16)
//     because a product of two fp16 values is exact in fp32 and fp32->fp16 double rounding of a two-term sum
//     cannot change the result.
#include <cuda_fp16.h>
#include <float.h>
#include <stdlib.h>

#include "matching_core.cuh"

namespace m3r {

#ifndef M3R_ITER_PROJ_TILE_DEFAULT
#define M3R_ITER_PROJ_TILE_DEFAULT 0
#endif
static int g_refine_mode = -1;  // -1: environment, 0: per-thread gathers, 1: tile-staged kernel where it applies

// matching_tile.cu
int launch_refine_tile_f16(const uint16_t* D11, const uint16_t* D21, const int64_t* p1, int64_t* p1_new, int b, int h,
                           int w, int fdim, int radius, int dilation_max, cudaStream_t st);

int launch_iter_proj_tile(const float4* P0, const float4* P1, const float* P2, const float* pts, const float* p_init,
                          float* p_new, uint8_t* converged, int b, int h, int w, int max_iter, int warm,
                          float lambda_init, float cost_thresh, cudaStream_t st);
static int g_iter_proj_mode = -1;  // -1: environment (M3R_ITER_PROJ_TILE=0/1, default below), 0: gathers, 1: tile-staged

// =================================================================================================== iter_proj
// 9-channel bilinear sample; weights follow the reference's "opposite area" naming (matching_kernels.cu:161-170):
//   w11 = du*dv multiplies pixel (v+1,u+1), w12 = (1-du)*dv -> (v+1,u), w21 = du*(1-dv) -> (v,u+1), w22 -> (v,u).
struct BilerpW {
  float w11, w12, w21, w22;
  int off22;  // element offset of pixel (v11,u11) channel 0 inside the image
};

template <bool kMirrorFp64>
__device__ __forceinline__ BilerpW bilerp_setup(float u, float v, int w) {
  BilerpW B;
  const int u11 = static_cast<int>(floorf(u));
  const int v11 = static_cast<int>(floorf(v));
  const float du = u - static_cast<float>(u11);
  const float dv = v - static_cast<float>(v11);
  B.w11 = du * dv;
  if (kMirrorFp64) {
    // the reference evaluates (1.0-du)*dv etc. in double and rounds once (matching_kernels.cu:162-164)
    B.w12 = (float)((1.0 - (double)du) * (double)dv);
    B.w21 = (float)((double)du * (1.0 - (double)dv));
    B.w22 = (float)((1.0 - (double)du) * (1.0 - (double)dv));
  } else {
    const float cu = 1.0f - du, cv = 1.0f - dv;
    B.w12 = cu * dv;
    B.w21 = du * cv;
    B.w22 = cu * cv;
  }
  B.off22 = (v11 * w + u11) * 9;
  return B;
}

template <bool kMirrorFp64>
__device__ __forceinline__ float inv_norm3(float x, float y, float z, float& nrm) {
  nrm = sqrtf(x * x + y * y + z * z);
  if (kMirrorFp64) return (float)(1.0 / (double)nrm);  // matching_kernels.cu:187
  return 1.0f / nrm;
}

template <bool kMirrorFp64>
__global__ void __launch_bounds__(128) iter_proj_kernel(const float* __restrict__ rays, const float* __restrict__ pts,
                                                        const float* __restrict__ p_init, float* __restrict__ p_new,
                                                        uint8_t* __restrict__ converged, long long total, int n, int h,
                                                        int w, int max_iter, float lambda_init, float cost_thresh) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int b = (int)(gid / n);
  const float* __restrict__ img = rays + (size_t)b * h * w * 9;
  const int row = w * 9;

  const float2 p0 = __ldg(reinterpret_cast<const float2*>(p_init) + gid);
  float u = fminf(fmaxf(p0.x, 1.0f), (float)(w - 2));
  float v = fminf(fmaxf(p0.y, 1.0f), (float)(h - 2));
  const float tx = __ldg(pts + gid * 3 + 0), ty = __ldg(pts + gid * 3 + 1), tz = __ldg(pts + gid * 3 + 2);

  float lambda = lambda_init;
  bool conv = false;
  for (int it = 0; it < max_iter; it++) {
    BilerpW B = bilerp_setup<kMirrorFp64>(u, v, w);
    const float* r22 = img + B.off22;  // (v, u)
    const float* r21 = r22 + 9;        // (v, u+1)
    const float* r12 = r22 + row;      // (v+1, u)
    const float* r11 = r12 + 9;        // (v+1, u+1)
    float s[9];
#pragma unroll
    for (int j = 0; j < 9; j++)
      s[j] = B.w11 * __ldg(r11 + j) + B.w12 * __ldg(r12 + j) + B.w21 * __ldg(r21 + j) + B.w22 * __ldg(r22 + j);
    float nrm;
    float inv = inv_norm3<kMirrorFp64>(s[0], s[1], s[2], nrm);
    const float e0 = s[0] * inv - tx, e1 = s[1] * inv - ty, e2 = s[2] * inv - tz;
    const float cost = e0 * e0 + e1 * e1 + e2 * e2;
    float A00 = s[3] * s[3] + s[4] * s[4] + s[5] * s[5];
    const float A01 = s[3] * s[6] + s[4] * s[7] + s[5] * s[8];
    float A11 = s[6] * s[6] + s[7] * s[7] + s[8] * s[8];
    const float b0 = -(e0 * s[3] + e1 * s[4] + e2 * s[5]);
    const float b1 = -(e0 * s[6] + e1 * s[7] + e2 * s[8]);
    A00 += lambda;
    A11 += lambda;
    const float det = A00 * A11 - A01 * A01;
    const float det_inv = kMirrorFp64 ? (float)(1.0 / (double)det) : 1.0f / det;
    const float delta_u = det_inv * (A11 * b0 - A01 * b1);
    const float delta_v = det_inv * (-A01 * b0 + A00 * b1);
    const float u_new = fminf(fmaxf(u + delta_u, 1.0f), (float)(w - 2));
    const float v_new = fminf(fmaxf(v + delta_v, 1.0f), (float)(h - 2));

    B = bilerp_setup<kMirrorFp64>(u_new, v_new, w);
    r22 = img + B.off22;
    r21 = r22 + 9;
    r12 = r22 + row;
    r11 = r12 + 9;
    float t[3];
#pragma unroll
    for (int j = 0; j < 3; j++)
      t[j] = B.w11 * __ldg(r11 + j) + B.w12 * __ldg(r12 + j) + B.w21 * __ldg(r21 + j) + B.w22 * __ldg(r22 + j);
    inv = inv_norm3<kMirrorFp64>(t[0], t[1], t[2], nrm);
    const float f0 = t[0] * inv - tx, f1 = t[1] * inv - ty, f2 = t[2] * inv - tz;
    const float new_cost = f0 * f0 + f1 * f1 + f2 * f2;
    if (new_cost < cost) {
      u = u_new;
      v = v_new;
      lambda = kMirrorFp64 ? (float)((double)lambda * 0.1) : lambda * 0.1f;
      conv = new_cost < cost_thresh;
    } else {
      lambda = kMirrorFp64 ? (float)((double)lambda * 10.0) : lambda * 10.0f;
      conv = cost < cost_thresh;
    }
  }
  reinterpret_cast<float2*>(p_new)[gid] = make_float2(u, v);
  converged[gid] = conv ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------ iter_proj v2
// Layout pre-pass + cached LM loop.  Profiling v1 on B200 showed the kernel bound by L1 wavefronts, not HBM or issue:
// every scalar LDG of the 36-byte-stride AoS image touches ~9 cache lines per warp and v1 issues 48 of them per LM
// iteration.  v2 repacks the 9-channel image once per call into 16-byte records
//   P0 = (ray.xyz, gx.x)   P1 = (gx.y, gx.z, gy.x, gy.y)   P2 = gy.z
// so a corner is 2 x LDG.128 + 1 x LDG.32, and keeps the last sample in registers: a rejected LM step re-uses the
// sample at (u,v) unchanged (the reference re-reads identical values), an accepted step re-uses the ray just sampled
// at the candidate position and only fetches the gradients there.
// Arithmetic notes (why this stays bit-compatible with matching_kernels.cu:155-268 up to rare double roundings):
//   * u,v >= 1 so du,dv are multiples of 2^-23 and 1-du, 1-dv are exact in fp32; the reference's double products
//     (1.0-du)*dv etc. rounded to float equal the single fp32 products;
//   * (float)(1.0/(double)x) equals the correctly rounded fp32 reciprocal unless the double quotient lands within
//     2^-54 of a float tie (probability ~2^-29 per operation);
//   * lambda*10.0 is exact-then-rounded in both precisions; lambda*0.1 is kept in double (0.1 != 0.1f).
__global__ void __launch_bounds__(256) repack_rays_kernel(const float* __restrict__ rays, float4* __restrict__ P0,
                                                          float4* __restrict__ P1, float* __restrict__ P2,
                                                          long long total) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= total) return;
  const float* s = rays + p * 9;
  P0[p] = make_float4(__ldg(s + 0), __ldg(s + 1), __ldg(s + 2), __ldg(s + 3));
  P1[p] = make_float4(__ldg(s + 4), __ldg(s + 5), __ldg(s + 6), __ldg(s + 7));
  P2[p] = __ldg(s + 8);
}

__global__ void __launch_bounds__(128) iter_proj_v2_kernel(const float4* __restrict__ P0, const float4* __restrict__ P1,
                                                           const float* __restrict__ P2, const float* __restrict__ pts,
                                                           const float* __restrict__ p_init, float* __restrict__ p_new,
                                                           uint8_t* __restrict__ converged, long long total, int n,
                                                           int h, int w, int max_iter, float lambda_init,
                                                           float cost_thresh) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int b = (int)(gid / n);
  const size_t img = (size_t)b * h * w;
  const float4* __restrict__ p0 = P0 + img;
  const float4* __restrict__ p1 = P1 + img;
  const float* __restrict__ p2 = P2 + img;

  const float2 pi = __ldg(reinterpret_cast<const float2*>(p_init) + gid);
  float u = fminf(fmaxf(pi.x, 1.0f), (float)(w - 2));
  float v = fminf(fmaxf(pi.y, 1.0f), (float)(h - 2));
  const float tx = __ldg(pts + gid * 3 + 0), ty = __ldg(pts + gid * 3 + 1), tz = __ldg(pts + gid * 3 + 2);

  bool conv = false;
  iter_proj_core(p0, p1, p2, tx, ty, tz, w, h, max_iter, lambda_init, cost_thresh, u, v, conv);
  reinterpret_cast<float2*>(p_new)[gid] = make_float2(u, v);
  converged[gid] = conv ? 1 : 0;
}

// ============================================================================================== refine_matches
// fp16, compile-time descriptor length (multiple of 8 halves = 16 bytes so rows are uint4-aligned).
template <int FDIM>
__global__ void __launch_bounds__(128) refine_matches_f16_kernel(const uint4* __restrict__ D11,
                                                                 const uint4* __restrict__ D21,
                                                                 const longlong2* __restrict__ p1,
                                                                 longlong2* __restrict__ p1_new, long long total,
                                                                 int n, int h, int w, int radius, int dilation_max) {
  constexpr int NV = FDIM / 8;  // uint4 per descriptor
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int b = (int)(gid / n);
  const longlong2 p = p1[gid];
  // coordinates far outside any image can never produce an in-image candidate: result == input
  const long long lim = 1LL << 29;
  if (p.x < -lim || p.x > lim || p.y < -lim || p.y > lim) {
    p1_new[gid] = p;
    return;
  }
  __half2 q[FDIM / 2];
  load_query_f16<FDIM>(D21, gid, q);
  int u_new, v_new;
  if (radius == 3)  // config/base.yaml matching.radius: the 7 candidates of a scan column are fetched together
    refine_core_f16<FDIM, 3>(q, D11 + (size_t)b * NV * h * w, NV, 1, h, w, 3, dilation_max, (int)p.x, (int)p.y, u_new,
                             v_new);
  else
    refine_core_f16<FDIM, 0>(q, D11 + (size_t)b * NV * h * w, NV, 1, h, w, radius, dilation_max, (int)p.x, (int)p.y,
                             u_new, v_new);
  p1_new[gid] = make_longlong2((long long)u_new, (long long)v_new);
}

// generic descriptor length (any fdim), scalar loads; T = __half (rounded like c10::Half) or float
template <typename T>
struct ScoreOps;
template <>
struct ScoreOps<__half> {
  __device__ static __forceinline__ __half zero() { return __ushort_as_half(0); }
  __device__ static __forceinline__ __half mac(__half s, __half a, __half b) {
    return __hadd_rn(s, __hmul_rn(a, b));
  }
  __device__ static __forceinline__ float to_float(__half s) { return __half2float(s); }
  __device__ static __forceinline__ float init_max() { return 0.0f; }
};
template <>
struct ScoreOps<float> {
  __device__ static __forceinline__ float zero() { return 0.0f; }
  __device__ static __forceinline__ float mac(float s, float a, float b) { return s + a * b; }
  __device__ static __forceinline__ float to_float(float s) { return s; }
  __device__ static __forceinline__ float init_max() { return FLT_MIN; }  // numeric_limits<float>::min()
};

template <typename T>
__global__ void __launch_bounds__(128) refine_matches_generic_kernel(const T* __restrict__ D11,
                                                                     const T* __restrict__ D21,
                                                                     const longlong2* __restrict__ p1,
                                                                     longlong2* __restrict__ p1_new, long long total,
                                                                     int n, int h, int w, int fdim, int radius,
                                                                     int dilation_max) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int b = (int)(gid / n);
  const longlong2 p = p1[gid];
  const long long lim = 1LL << 29;
  if (p.x < -lim || p.x > lim || p.y < -lim || p.y > lim) {
    p1_new[gid] = p;
    return;
  }
  const T* __restrict__ qd = D21 + (size_t)gid * fdim;
  const T* __restrict__ img = D11 + (size_t)b * h * w * fdim;
  int u0 = (int)p.x, v0 = (int)p.y;
  int u_new = u0, v_new = v0;
  float max_score = ScoreOps<T>::init_max();
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    const int diam = 2 * rd + 1;
    for (int i = 0; i < diam; i += d) {
      const int u = u0 - rd + i;
      if (u < 0 || u >= w) continue;
      for (int j = 0; j < diam; j += d) {
        const int v = v0 - rd + j;
        if (v < 0 || v >= h) continue;
        const T* c = img + ((size_t)v * w + u) * fdim;
        T score = ScoreOps<T>::zero();
        for (int k = 0; k < fdim; k++) score = ScoreOps<T>::mac(score, __ldg(qd + k), __ldg(c + k));
        const float s = ScoreOps<T>::to_float(score);
        if (s > max_score) {
          max_score = s;
          u_new = u;
          v_new = v;
        }
      }
    }
    u0 = u_new;
    v0 = v_new;
  }
  p1_new[gid] = make_longlong2((long long)u_new, (long long)v_new);
}

}  // namespace m3r

// ======================================================================================================= C ABI
using namespace m3r;

extern "C" int m3r_iter_proj(const float* rays_img_with_grad, const float* pts_3d_norm, const float* p_init,
                             float* p_new, uint8_t* converged, int b, int h, int w, int n, int max_iter,
                             float lambda_init, float cost_thresh, void* stream) {
  M3R_REQUIRE(b >= 0 && n >= 0 && h >= 0 && w >= 0 && max_iter >= 0, "iter_proj: negative size");
  const long long total = (long long)b * n;
  if (total == 0) return M3R_OK;
  M3R_REQUIRE(rays_img_with_grad && pts_3d_norm && p_init && p_new && converged, "iter_proj: null pointer");
  M3R_REQUIRE(h >= 3 && w >= 3, "iter_proj: image must be at least 3x3 (got %dx%d)", h, w);
  M3R_REQUIRE((long long)h * w * 9 < (1LL << 31), "iter_proj: image too large for 32-bit offsets");
  M3R_REQUIRE(((uintptr_t)p_init & 7) == 0 && ((uintptr_t)p_new & 7) == 0, "iter_proj: p_init/p_new must be 8-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int threads = 128;
  const unsigned blocks = (unsigned)ceil_div_ll(total, threads);
  static const bool use_v1 = getenv("M3R_ITER_PROJ_V1") != nullptr;  // A/B switch for profiling only
  if (use_v1) {
    iter_proj_kernel<true><<<blocks, threads, 0, st>>>(rays_img_with_grad, pts_3d_norm, p_init, p_new, converged, total,
                                                       n, h, w, max_iter, lambda_init, cost_thresh);
    M3R_LAUNCH_CHECK();
    return M3R_OK;
  }
  // stream-ordered scratch for the repacked image (36 B per pixel, 16-byte aligned planes); no host sync
  const long long npix = (long long)b * h * w;
  const size_t plane = align_up((size_t)npix * sizeof(float4), 256);
  {
    const int rc = keep_async_pool();
    if (rc != M3R_OK) return rc;
  }
  void* scratch = nullptr;
  M3R_CUDA(cudaMallocAsync(&scratch, 2 * plane + align_up((size_t)npix * sizeof(float), 256), st));
  float4* P0 = (float4*)scratch;
  float4* P1 = (float4*)((char*)scratch + plane);
  float* P2 = (float*)((char*)scratch + 2 * plane);
  repack_rays_kernel<<<(unsigned)ceil_div_ll(npix, 256), 256, 0, st>>>(rays_img_with_grad, P0, P1, P2, npix);
  M3R_LAUNCH_CHECK();
  // queries on the image grid (the callers' layout): the tile-staged kernel (matching_tile.cu); anything else, or a
  // shape it does not take: per-thread gathers
  static const int env_tile = [] {
    const char* e = getenv("M3R_ITER_PROJ_TILE");
    return e == nullptr ? M3R_ITER_PROJ_TILE_DEFAULT : atoi(e);
  }();
  static const int env_warm = [] {
    const char* e = getenv("M3R_ITER_PROJ_WARM");
    return e == nullptr ? 1 : atoi(e);
  }();
  int rc = 1;
  if ((g_iter_proj_mode == 1 || (g_iter_proj_mode < 0 && env_tile != 0)) && (long long)n == (long long)h * w)
    rc = launch_iter_proj_tile(P0, P1, P2, pts_3d_norm, p_init, p_new, converged, b, h, w, max_iter, env_warm,
                               lambda_init, cost_thresh, st);
  if (rc == 1) {
    iter_proj_v2_kernel<<<blocks, threads, 0, st>>>(P0, P1, P2, pts_3d_norm, p_init, p_new, converged, total, n, h, w,
                                                    max_iter, lambda_init, cost_thresh);
    M3R_LAUNCH_CHECK();
    rc = M3R_OK;
  }
  M3R_CUDA(cudaFreeAsync(scratch, st));
  return rc;
}

extern "C" void m3r_debug_set_iter_proj_mode(int mode) { m3r::g_iter_proj_mode = mode; }

// ---------------------------------------------------------------------------------------------------- no-allocation variant
// m3r_iter_proj for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc): the scratch of the repacked ray image (three planes, 36 bytes per pixel) lives in a file-scope device
// array of fixed size instead of stream-ordered memory.  Every GPU that loads the library holds these 6.75 MiB until the
// process exits.  In host code the array's NAME is only a host-side stand-in: its device address is looked up once per
// device with cudaGetSymbolAddress (that device current) and handed to the kernels.
static __device__ __align__(256) unsigned char g_iter_proj_scratch[M3R_ITER_PROJ_NOALLOC_SCRATCH_BYTES];

extern "C" int m3r_iter_proj_noalloc(const float* rays_img_with_grad, const float* pts_3d_norm, const float* p_init,
                                     float* p_new, uint8_t* converged, int b, int h, int w, int n, int max_iter,
                                     float lambda_init, float cost_thresh, void* stream) {
  M3R_REQUIRE(b >= 0 && n >= 0 && h >= 0 && w >= 0 && max_iter >= 0, "iter_proj_noalloc: negative size");
  const long long total = (long long)b * n;
  if (total == 0) return M3R_OK;
  M3R_REQUIRE(rays_img_with_grad && pts_3d_norm && p_init && p_new && converged, "iter_proj_noalloc: null pointer");
  M3R_REQUIRE(h >= 3 && w >= 3, "iter_proj_noalloc: image must be at least 3x3 (got %dx%d)", h, w);
  M3R_REQUIRE(((uintptr_t)p_init & 7) == 0 && ((uintptr_t)p_new & 7) == 0,
              "iter_proj_noalloc: p_init/p_new must be 8-byte aligned");
  const long long npix = (long long)b * h * w;
  M3R_REQUIRE(npix <= M3R_ITER_PROJ_NOALLOC_MAX_PIXELS,
              "iter_proj_noalloc: b*h*w = %lld pixels of ray image exceed the fixed scratch (at most %d pixels, e.g. one "
              "512x384 image); use m3r_iter_proj for larger batches", npix, M3R_ITER_PROJ_NOALLOC_MAX_PIXELS);
  // device address of the scratch, once per device (cached: the look-up itself allocates nothing)
  static unsigned char* scratch_dev[kMaxDevices] = {};
  const int slot = current_device_slot();
  if (scratch_dev[slot] == nullptr) {
    void* ptr = nullptr;
    M3R_CUDA(cudaGetSymbolAddress(&ptr, g_iter_proj_scratch));
    M3R_REQUIRE(ptr != nullptr, "iter_proj_noalloc: cudaGetSymbolAddress returned a null scratch address");
    scratch_dev[slot] = (unsigned char*)ptr;
  }
  // plane offsets for the MAXIMUM size, so they do not depend on this call's npix (16-byte aligned by construction)
  unsigned char* base = scratch_dev[slot];
  float4* P0 = (float4*)base;
  float4* P1 = (float4*)(base + (size_t)M3R_ITER_PROJ_NOALLOC_MAX_PIXELS * sizeof(float4));
  float* P2 = (float*)(base + 2 * (size_t)M3R_ITER_PROJ_NOALLOC_MAX_PIXELS * sizeof(float4));
  cudaStream_t st = (cudaStream_t)stream;
  // the repack writes every plane entry [0, npix) that the LM kernel can read, so nothing of an earlier call's contents is
  // ever consumed and nothing needs clearing
  repack_rays_kernel<<<(unsigned)ceil_div_ll(npix, 256), 256, 0, st>>>(rays_img_with_grad, P0, P1, P2, npix);
  M3R_LAUNCH_CHECK();
  iter_proj_v2_kernel<<<(unsigned)ceil_div_ll(total, 128), 128, 0, st>>>(P0, P1, P2, pts_3d_norm, p_init, p_new, converged,
                                                                         total, n, h, w, max_iter, lambda_init, cost_thresh);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

template <typename T>
static int launch_refine_generic(const T* D11, const T* D21, const int64_t* p1, int64_t* p1_new, long long total, int n,
                                 int h, int w, int fdim, int radius, int dilation_max, cudaStream_t st) {
  const int threads = 128;
  const unsigned blocks = (unsigned)ceil_div_ll(total, threads);
  refine_matches_generic_kernel<T><<<blocks, threads, 0, st>>>(D11, D21, (const longlong2*)p1, (longlong2*)p1_new, total,
                                                               n, h, w, fdim, radius, dilation_max);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

static int refine_check(const void* D11, const void* D21, const void* p1, const void* p1_new, int b, int h, int w,
                        int n, int fdim, int radius, int dilation_max) {
  M3R_REQUIRE(b >= 0 && n >= 0 && h >= 0 && w >= 0 && fdim >= 0, "refine_matches: negative size");
  if ((long long)b * n == 0) return 1;
  M3R_REQUIRE(D11 && D21 && p1 && p1_new, "refine_matches: null pointer");
  M3R_REQUIRE(radius >= 0 && dilation_max >= 0, "refine_matches: radius/dilation must be >= 0");
  M3R_REQUIRE(((uintptr_t)p1 & 15) == 0 && ((uintptr_t)p1_new & 15) == 0, "refine_matches: p1/p1_new must be 16-byte aligned");
  M3R_REQUIRE((long long)radius * dilation_max < (1 << 20), "refine_matches: window too large");
  return 0;
}

extern "C" int m3r_refine_matches_f16(const uint16_t* D11, const uint16_t* D21, const int64_t* p1, int64_t* p1_new,
                                      int b, int h, int w, int n, int fdim, int radius, int dilation_max,
                                      void* stream) {
  int c = refine_check(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max);
  if (c != 0) return c < 0 ? c : M3R_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const long long total = (long long)b * n;
  const bool aligned = (((uintptr_t)D11 | (uintptr_t)D21) & 15) == 0;
  if ((fdim == 24 || fdim == 16) && aligned && (long long)h * w * (fdim / 8) < (1LL << 31)) {
    // queries on the pixel grid of an image (every caller of the reference, matching.py:79-85): the tile-staged kernel of
    // matching_tile.cu is available as an opt-in (m3r_debug_set_refine_mode(1) or M3R_REFINE_TILE=1).  Measured r2 at
    // 512x384, b = 1: 0.302 ms against 0.338 ms for the per-thread gathers on the noisy scene, 0.209 against 0.158 ms on
    // the noise-free one, where neighbouring queries already share L1 lines and the tile kernel's 2 CTAs per SM cost more
    // than its shared-memory reads save -- so the gathers stay the default.
    static const bool env_tile = getenv("M3R_REFINE_TILE") != nullptr;
    if ((g_refine_mode == 1 || (g_refine_mode < 0 && env_tile)) && (long long)n == (long long)h * w) {
      const int rc = launch_refine_tile_f16(D11, D21, p1, p1_new, b, h, w, fdim, radius, dilation_max, st);
      if (rc != 1) return rc;  // 1: shape / parameters outside the tile kernel's range -> per-thread gathers below
    }
    const int threads = 128;
    const unsigned blocks = (unsigned)ceil_div_ll(total, threads);
    if (fdim == 24)
      refine_matches_f16_kernel<24><<<blocks, threads, 0, st>>>((const uint4*)D11, (const uint4*)D21,
                                                                (const longlong2*)p1, (longlong2*)p1_new, total, n, h, w,
                                                                radius, dilation_max);
    else
      refine_matches_f16_kernel<16><<<blocks, threads, 0, st>>>((const uint4*)D11, (const uint4*)D21,
                                                                (const longlong2*)p1, (longlong2*)p1_new, total, n, h, w,
                                                                radius, dilation_max);
    M3R_LAUNCH_CHECK();
    return M3R_OK;
  }
  return launch_refine_generic<__half>((const __half*)D11, (const __half*)D21, p1, p1_new, total, n, h, w, fdim, radius,
                                       dilation_max, st);
}

extern "C" void m3r_debug_set_refine_mode(int mode) { m3r::g_refine_mode = mode; }

extern "C" int m3r_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1, int64_t* p1_new, int b,
                                      int h, int w, int n, int fdim, int radius, int dilation_max, void* stream) {
  int c = refine_check(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max);
  if (c != 0) return c < 0 ? c : M3R_OK;
  return launch_refine_generic<float>(D11, D21, p1, p1_new, (long long)b * n, n, h, w, fdim, radius, dilation_max,
                                      (cudaStream_t)stream);
}
