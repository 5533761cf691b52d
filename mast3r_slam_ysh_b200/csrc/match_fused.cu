// Fused dense matching for sm_100a: the whole of match_iterative_proj (mast3r_slam/matching.py:52-91) on the device
// in five launches -- SURVEY.md section 8(f) rank 1 ("fuse the Python pre/post of matching").  The reference builds a
// 9-channel ray image with ~6 eager torch ops (normalize, two depthwise 3x3 convolutions on a reflect-padded copy, cat,
// permute, contiguous; matching.py:25-49, image.py:5-38), normalises the query points, expands the initial indices to
// float pixels, calls iter_proj, gathers X11 at the matches for the occlusion test (matching.py:70-76), calls
// refine_matches and converts pixels to linear indices (:88).  Here:
//   ray_normalize_kernel : X11 -> unit rays (float4)
//   ray_planes_kernel    : rays -> the three 16-byte planes iter_proj samples (ray + Scharr d/du, d/dv, reflect border);
//                          the 36-byte AoS image and its repack are never materialised
//   match_proj_kernel    : normalise X21 on the fly, initial pixel from the index, 10 LM iterations (iter_proj_core),
//                          truncation to integer pixels, occlusion test against X11, valid flag
//   desc_half_kernel     : D11 fp32 -> fp16 (the .half() of matching.py:80)
//   match_refine_kernel  : D21 -> fp16 on the fly, descriptor refinement (refine_core_f16) + pixel -> linear index
// Additive entry point (m3r_match); the five legacy signatures are untouched.
#include "matching_core.cuh"

namespace m3r {

__global__ void __launch_bounds__(256) ray_normalize_kernel(const float* __restrict__ X, float4* __restrict__ R,
                                                            long long npix) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npix) return;
  const float x = __ldg(X + 3 * p), y = __ldg(X + 3 * p + 1), z = __ldg(X + 3 * p + 2);
  // F.normalize(dim=-1): x / max(||x||_2, 1e-12)   (matching.py:30)
  const float nrm = fmaxf(sqrtf(x * x + y * y + z * z), 1e-12f);
  R[p] = make_float4(__fdiv_rn(x, nrm), __fdiv_rn(y, nrm), __fdiv_rn(z, nrm), 0.0f);
}

__device__ __forceinline__ int reflect(int i, int n) { return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i); }

// img_gradient (image.py:5-38): depthwise 3x3 cross-correlation with (1/32)[[-3,0,3],[-10,0,10],[-3,0,3]] (d/du) and its
// transpose (d/dv) on the reflect-padded ray image.
template <bool AOS>
__global__ void __launch_bounds__(256) ray_planes_kernel(const float4* __restrict__ R, float4* __restrict__ P0,
                                                         float4* __restrict__ P1, float* __restrict__ P2,
                                                         float* __restrict__ aos, int h, int w, long long npix) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npix) return;
  const long long img = p / ((long long)h * w);
  const int rem = (int)(p - img * h * w);
  const int v = rem / w, u = rem - v * w;
  const float4* __restrict__ I = R + img * h * w;
  const int um = reflect(u - 1, w), up = reflect(u + 1, w), vm = reflect(v - 1, h), vp = reflect(v + 1, h);
  const float4 a = __ldg(I + vm * w + um), b = __ldg(I + vm * w + u), c = __ldg(I + vm * w + up);
  const float4 d = __ldg(I + v * w + um), e = __ldg(I + v * w + u), f = __ldg(I + v * w + up);
  const float4 g = __ldg(I + vp * w + um), hh = __ldg(I + vp * w + u), i = __ldg(I + vp * w + up);
  const float k3 = 3.0f / 32.0f, k10 = 10.0f / 32.0f;
  const float gx0 = k3 * (c.x - a.x) + k10 * (f.x - d.x) + k3 * (i.x - g.x);
  const float gx1 = k3 * (c.y - a.y) + k10 * (f.y - d.y) + k3 * (i.y - g.y);
  const float gx2 = k3 * (c.z - a.z) + k10 * (f.z - d.z) + k3 * (i.z - g.z);
  const float gy0 = k3 * (g.x - a.x) + k10 * (hh.x - b.x) + k3 * (i.x - c.x);
  const float gy1 = k3 * (g.y - a.y) + k10 * (hh.y - b.y) + k3 * (i.y - c.y);
  const float gy2 = k3 * (g.z - a.z) + k10 * (hh.z - b.z) + k3 * (i.z - c.z);
  if (AOS) {  // the reference's (b,h,w,9) layout: ray, d/du, d/dv (matching.py:33-36)
    float* o = aos + p * 9;
    o[0] = e.x; o[1] = e.y; o[2] = e.z; o[3] = gx0; o[4] = gx1; o[5] = gx2; o[6] = gy0; o[7] = gy1; o[8] = gy2;
  } else {
    P0[p] = make_float4(e.x, e.y, e.z, gx0);
    P1[p] = make_float4(gx1, gx2, gy0, gy1);
    P2[p] = gy2;
  }
}

__global__ void __launch_bounds__(128) match_proj_kernel(const float4* __restrict__ P0, const float4* __restrict__ P1,
                                                         const float* __restrict__ P2, const float* __restrict__ X11,
                                                         const float* __restrict__ X21,
                                                         const int64_t* __restrict__ idx_init, int2* __restrict__ p1_out,
                                                         uint8_t* __restrict__ valid, long long total, int n, int h,
                                                         int w, int max_iter, float lambda_init, float cost_thresh,
                                                         float dist_thresh) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int b = (int)(gid / n);
  const int k = (int)(gid - (long long)b * n);
  const size_t img = (size_t)b * h * w;
  // query ray: F.normalize(X21) (matching.py:40-41)
  const float qx = __ldg(X21 + gid * 3), qy = __ldg(X21 + gid * 3 + 1), qz = __ldg(X21 + gid * 3 + 2);
  const float nrm = fmaxf(sqrtf(qx * qx + qy * qy + qz * qz), 1e-12f);
  const float tx = __fdiv_rn(qx, nrm), ty = __fdiv_rn(qy, nrm), tz = __fdiv_rn(qz, nrm);
  // initial pixel: lin_to_pixel(idx) or the identity mapping (matching.py:44-48, :19-23)
  long long lin = idx_init != nullptr ? (long long)idx_init[gid] : (long long)k;
  float u = (float)(lin % w), v = (float)(lin / w);
  u = fminf(fmaxf(u, 1.0f), (float)(w - 2));
  v = fminf(fmaxf(v, 1.0f), (float)(h - 2));
  bool conv = false;
  iter_proj_core(P0 + img, P1 + img, P2 + img, tx, ty, tz, w, h, max_iter, lambda_init, cost_thresh, u, v, conv);
  const int pu = (int)u, pv = (int)v;  // p1.long(): truncation (matching.py:68); u, v >= 1
  // occlusion test (matching.py:70-76): || X11[b, pv, pu] - X21[b, k] || < dist_thresh
  const float* a = X11 + (img + (size_t)pv * w + pu) * 3;
  const float dx = __ldg(a) - qx, dy = __ldg(a + 1) - qy, dz = __ldg(a + 2) - qz;
  const float dist = sqrtf(dx * dx + dy * dy + dz * dz);
  p1_out[gid] = make_int2(pu, pv);
  valid[gid] = (conv && dist < dist_thresh) ? 1 : 0;
}

// fp32 (h,w,fdim) descriptors -> fp16, same layout (the .half() of matching.py:80).  A chunk-planar repack was tried
// and measured no faster (0.377 vs 0.386 ms at 512x384): the refinement is bound by L1 tag look-ups of scattered
// 16-byte gathers, not by the 48-byte stride (tools/time_refine.py, r1).
__global__ void __launch_bounds__(256) desc_half_kernel(const float* __restrict__ D, uint4* __restrict__ P,
                                                        long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one 8-float chunk per thread
  if (t >= total) return;
  const float4 a = __ldg(reinterpret_cast<const float4*>(D) + 2 * t), c = __ldg(reinterpret_cast<const float4*>(D) + 2 * t + 1);
  uint4 o;
  *reinterpret_cast<__half2*>(&o.x) = __floats2half2_rn(a.x, a.y);
  *reinterpret_cast<__half2*>(&o.y) = __floats2half2_rn(a.z, a.w);
  *reinterpret_cast<__half2*>(&o.z) = __floats2half2_rn(c.x, c.y);
  *reinterpret_cast<__half2*>(&o.w) = __floats2half2_rn(c.z, c.w);
  P[t] = o;
}

template <int FDIM>
__global__ void __launch_bounds__(128) match_refine_kernel(const uint4* __restrict__ planes, const float* __restrict__ D21,
                                                           const int2* __restrict__ p1, int64_t* __restrict__ idx_out,
                                                           long long total, int n, int h, int w, int radius,
                                                           int dilation_max) {
  constexpr int NV = FDIM / 8;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int b = (int)(gid / n);
  const int2 p = p1[gid];
  __half2 q[FDIM / 2];
  load_query_f32<FDIM>(D21, gid, q);
  int u_new, v_new;
  if (radius == 3)
    refine_core_f16<FDIM, 3>(q, planes + (size_t)b * NV * h * w, NV, 1, h, w, 3, dilation_max, p.x, p.y, u_new, v_new);
  else
    refine_core_f16<FDIM, 0>(q, planes + (size_t)b * NV * h * w, NV, 1, h, w, radius, dilation_max, p.x, p.y, u_new,
                             v_new);
  idx_out[gid] = (int64_t)u_new + (int64_t)w * (int64_t)v_new;  // pixel_to_lin (matching.py:14-16)
}

__global__ void __launch_bounds__(256) match_lin_kernel(const int2* __restrict__ p1, int64_t* __restrict__ idx_out,
                                                        long long total, int w) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int2 p = p1[gid];
  idx_out[gid] = (int64_t)p.x + (int64_t)w * (int64_t)p.y;
}

}  // namespace m3r

using namespace m3r;

extern "C" int m3r_match(const float* X11, const float* X21, const float* D11, const float* D21,
                         const int64_t* idx_1_to_2_init, int64_t* idx_1_to_2, uint8_t* valid_match, int b, int h, int w,
                         int fdim, int max_iter, float lambda_init, float cost_thresh, float dist_thresh, int radius,
                         int dilation_max, void* stream) {
  M3R_REQUIRE(b >= 0 && h >= 0 && w >= 0 && max_iter >= 0 && radius >= 0 && dilation_max >= 0, "match: negative size");
  const long long npix = (long long)b * h * w;
  if (npix == 0) return M3R_OK;
  M3R_REQUIRE(X11 && X21 && idx_1_to_2 && valid_match, "match: null pointer");
  M3R_REQUIRE(h >= 3 && w >= 3, "match: image must be at least 3x3 (got %dx%d)", h, w);
  M3R_REQUIRE((long long)h * w * 9 < (1LL << 31), "match: image too large for 32-bit offsets");
  const bool refine = radius > 0 && dilation_max > 0;
  if (refine) {
    M3R_REQUIRE(D11 && D21, "match: null descriptor pointer");
    if (!((fdim == 24 || fdim == 16) && ((((uintptr_t)D11 | (uintptr_t)D21) & 15) == 0))) {
      set_error("match: fused path needs 16-byte aligned fp32 descriptors with fdim 16 or 24 (got %d)", fdim);
      return M3R_ERR_UNSUPPORTED;
    }
  }
  cudaStream_t st = (cudaStream_t)stream;
  const size_t plane = align_up((size_t)npix * sizeof(float4), 256);
  const size_t p2b = align_up((size_t)npix * sizeof(float), 256);
  const size_t p1b = align_up((size_t)npix * sizeof(int2), 256);
  const int nv = refine ? fdim / 8 : 0;
  const size_t dpb = align_up((size_t)npix * nv * sizeof(uint4), 256);
  {
    const int rc = keep_async_pool();
    if (rc != M3R_OK) return rc;
  }
  void* scratch = nullptr;
  M3R_CUDA(cudaMallocAsync(&scratch, 3 * plane + p2b + p1b + dpb, st));
  float4* R = (float4*)scratch;
  float4* P0 = (float4*)((char*)scratch + plane);
  float4* P1 = (float4*)((char*)scratch + 2 * plane);
  float* P2 = (float*)((char*)scratch + 3 * plane);
  int2* p1 = (int2*)((char*)scratch + 3 * plane + p2b);
  uint4* dplanes = (uint4*)((char*)scratch + 3 * plane + p2b + p1b);
  const unsigned g256 = (unsigned)ceil_div_ll(npix, 256), g128 = (unsigned)ceil_div_ll(npix, 128);
  ray_normalize_kernel<<<g256, 256, 0, st>>>(X11, R, npix);
  M3R_LAUNCH_CHECK();
  ray_planes_kernel<false><<<g256, 256, 0, st>>>(R, P0, P1, P2, nullptr, h, w, npix);
  M3R_LAUNCH_CHECK();
  match_proj_kernel<<<g128, 128, 0, st>>>(P0, P1, P2, X11, X21, idx_1_to_2_init, p1, valid_match, npix, h * w, h, w,
                                          max_iter, lambda_init, cost_thresh, dist_thresh);
  M3R_LAUNCH_CHECK();
  if (!refine) {
    match_lin_kernel<<<g256, 256, 0, st>>>(p1, idx_1_to_2, npix, w);
  } else {
    const long long chunks = npix * nv;
    desc_half_kernel<<<(unsigned)ceil_div_ll(chunks, 256), 256, 0, st>>>(D11, dplanes, chunks);
    M3R_LAUNCH_CHECK();
    if (fdim == 24)
      match_refine_kernel<24><<<g128, 128, 0, st>>>(dplanes, D21, p1, idx_1_to_2, npix, h * w, h, w, radius, dilation_max);
    else
      match_refine_kernel<16><<<g128, 128, 0, st>>>(dplanes, D21, p1, idx_1_to_2, npix, h * w, h, w, radius, dilation_max);
  }
  M3R_LAUNCH_CHECK();
  M3R_CUDA(cudaFreeAsync(scratch, st));
  return M3R_OK;
}

extern "C" int m3r_ray_image(const float* X11, float* rays_img_with_grad, int b, int h, int w, void* stream) {
  M3R_REQUIRE(b >= 0 && h >= 0 && w >= 0, "ray_image: negative size");
  const long long npix = (long long)b * h * w;
  if (npix == 0) return M3R_OK;
  M3R_REQUIRE(X11 && rays_img_with_grad, "ray_image: null pointer");
  M3R_REQUIRE(h >= 2 && w >= 2, "ray_image: reflect padding needs at least 2x2 (got %dx%d)", h, w);
  cudaStream_t st = (cudaStream_t)stream;
  {
    const int rc = keep_async_pool();
    if (rc != M3R_OK) return rc;
  }
  void* scratch = nullptr;
  M3R_CUDA(cudaMallocAsync(&scratch, (size_t)npix * sizeof(float4), st));
  const unsigned g256 = (unsigned)ceil_div_ll(npix, 256);
  ray_normalize_kernel<<<g256, 256, 0, st>>>(X11, (float4*)scratch, npix);
  M3R_LAUNCH_CHECK();
  ray_planes_kernel<true><<<g256, 256, 0, st>>>((const float4*)scratch, nullptr, nullptr, nullptr, rays_img_with_grad, h, w,
                                                npix);
  M3R_LAUNCH_CHECK();
  M3R_CUDA(cudaFreeAsync(scratch, st));
  return M3R_OK;
}
