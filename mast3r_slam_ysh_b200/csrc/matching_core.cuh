// Device-side cores of the matching kernels, shared by matching.cu (drop-in entry points) and match_fused.cu.
#pragma once
#include <cuda_fp16.h>

#include "m3r_common.cuh"

namespace m3r {

struct Bil {
  float w11, w12, w21, w22;
  int u11, v11;  // the cell: its corners are pixels (v,u) "22", (v,u+1) "21", (v+1,u) "12", (v+1,u+1) "11"
};

__device__ __forceinline__ Bil bil_setup(float u, float v, int w) {
  Bil B;
  const int u11 = static_cast<int>(floorf(u));
  const int v11 = static_cast<int>(floorf(v));
  const float du = __fsub_rn(u, static_cast<float>(u11));
  const float dv = __fsub_rn(v, static_cast<float>(v11));
  const float cu = __fsub_rn(1.0f, du), cv = __fsub_rn(1.0f, dv);  // exact (see header comment)
  B.w11 = __fmul_rn(du, dv);
  B.w12 = __fmul_rn(cu, dv);
  B.w21 = __fmul_rn(du, cv);
  B.w22 = __fmul_rn(cu, cv);
  B.u11 = u11;
  B.v11 = v11;
  return B;
}

// The reference binary (nvcc 12.9, default -fmad=true) evaluates its float expressions with these exact fused
// sequences (read off its SASS); they are spelled out with non-contracting intrinsics so that this kernel's
// rounding does not depend on how the compiler happens to contract a differently shaped program.
//   w11*a11 + w12*a12 + w21*a21 + w22*a22  ->  fma(w22,a22, fma(w21,a21, fma(w11,a11, w12*a12)))
//   x*x' + y*y' + z*z'                     ->  fma(z,z', fma(x,x', y*y'))
__device__ __forceinline__ float bil(const Bil& B, float a11, float a12, float a21, float a22) {
  return __fmaf_rn(B.w22, a22, __fmaf_rn(B.w21, a21, __fmaf_rn(B.w11, a11, __fmul_rn(B.w12, a12))));
}
__device__ __forceinline__ float dot3_ref(float x, float xp, float y, float yp, float z, float zp) {
  return __fmaf_rn(z, zp, __fmaf_rn(x, xp, __fmul_rn(y, yp)));
}

// One query of iter_proj (matching_kernels.cu:119-275) on the repacked planes: 10 LM iterations from the clamped (u, v)
// towards the unit target ray (tx, ty, tz).  Shared by the drop-in kernel and the fused match kernel.
// Tried and dropped (r2, tools/time_iterproj.py, 512x384): keeping the four corners of the current bilinear cell in
// registers, tagged with the cell index, so that a candidate in the same cell (70-95 % of them from the third iteration on)
// needs no loads.  Bit-identical, but SLOWER at b >= 8 -- noise-free scene 1.07 -> 1.58 ms at b = 64, noisy scene 2.39 ->
// 2.59 ms (b = 1: 0.049 -> 0.047 ms): the repeated loads were L1 hits anyway, and 104 instead of 62 registers plus the
// divergent "same cell?" paths cost more than they saved.

// Where the corner samples come from.  ray(): plane P0 at the four corners of B's cell; grad(): planes P1 and P2.
struct GlobalRaySampler {
  const float4* __restrict__ p0;
  const float4* __restrict__ p1;
  const float* __restrict__ p2;
  int w;
  // one address per image row of the cell, the second corner of a row at an immediate offset (the kernel is issue-bound
  // on smooth input, profiles/r2_iterproj_b64_ncu_full.txt, and four independent 64-bit addresses per fetch were a
  // noticeable share of its instructions)
  __device__ __forceinline__ void ray(const Bil& B, float4& a11, float4& a12, float4& a21, float4& a22) const {
    const float4* __restrict__ r0 = p0 + (B.v11 * w + B.u11);
    const float4* __restrict__ r1 = r0 + w;
    a22 = __ldg(r0);
    a21 = __ldg(r0 + 1);
    a12 = __ldg(r1);
    a11 = __ldg(r1 + 1);
  }
  __device__ __forceinline__ void grad(const Bil& B, float4& c11, float4& c12, float4& c21, float4& c22, float& d11,
                                       float& d12, float& d21, float& d22) const {
    const int o = B.v11 * w + B.u11;
    const float4* __restrict__ r0 = p1 + o;
    const float4* __restrict__ r1 = r0 + w;
    c22 = __ldg(r0);
    c21 = __ldg(r0 + 1);
    c12 = __ldg(r1);
    c11 = __ldg(r1 + 1);
    const float* __restrict__ q0 = p2 + o;
    const float* __restrict__ q1 = q0 + w;
    d22 = __ldg(q0);
    d21 = __ldg(q0 + 1);
    d12 = __ldg(q1);
    d11 = __ldg(q1 + 1);
  }
};

// Everything one query carries from one LM iteration to the next, so that the loop can be run in pieces (the tile-staged
// kernel changes its sampler after the first iteration).
struct IterProjState {
  float u, v, lambda;
  bool conv, have_r, have_g;
  // sample cached at (u,v): raw ray, d/du, d/dv
  float r0, r1, r2, gx0, gx1, gx2, gy0, gy1, gy2;
  Bil B;
};

__device__ __forceinline__ void iter_proj_begin(IterProjState& s, float u, float v, int w, float lambda_init) {
  s.u = u;
  s.v = v;
  s.lambda = lambda_init;
  s.conv = false;
  s.have_r = false;
  s.have_g = false;
  s.r0 = s.r1 = s.r2 = s.gx0 = s.gx1 = s.gx2 = s.gy0 = s.gy1 = s.gy2 = 0.f;
  s.B = bil_setup(u, v, w);
}

// LM iterations [it0, it1) of one query (matching_kernels.cu:155-268)
template <typename Sampler>
__device__ __forceinline__ void iter_proj_steps(IterProjState& s, const Sampler& S, float tx, float ty, float tz, int w,
                                                int h, int it0, int it1, float cost_thresh) {
  for (int it = it0; it < it1; it++) {
    if (!s.have_r) {
      float4 a11, a12, a21, a22;
      S.ray(s.B, a11, a12, a21, a22);
      s.r0 = bil(s.B, a11.x, a12.x, a21.x, a22.x);
      s.r1 = bil(s.B, a11.y, a12.y, a21.y, a22.y);
      s.r2 = bil(s.B, a11.z, a12.z, a21.z, a22.z);
      s.gx0 = bil(s.B, a11.w, a12.w, a21.w, a22.w);
    }
    if (!s.have_g) {
      float4 c11, c12, c21, c22;
      float d11, d12, d21, d22;
      S.grad(s.B, c11, c12, c21, c22, d11, d12, d21, d22);
      s.gx1 = bil(s.B, c11.x, c12.x, c21.x, c22.x);
      s.gx2 = bil(s.B, c11.y, c12.y, c21.y, c22.y);
      s.gy0 = bil(s.B, c11.z, c12.z, c21.z, c22.z);
      s.gy1 = bil(s.B, c11.w, c12.w, c21.w, c22.w);
      s.gy2 = bil(s.B, d11, d12, d21, d22);
    }
    const float r0 = s.r0, r1 = s.r1, r2 = s.r2, gx0 = s.gx0, gx1 = s.gx1, gx2 = s.gx2, gy0 = s.gy0, gy1 = s.gy1,
                gy2 = s.gy2;
    // normalise, residual, cost (matching_kernels.cu:186-198); 1.0/r_norm compiles to the IEEE fp32 reciprocal
    float inv = __frcp_rn(__fsqrt_rn(dot3_ref(r0, r0, r1, r1, r2, r2)));
    const float e0 = __fmaf_rn(r0, inv, -tx), e1 = __fmaf_rn(r1, inv, -ty), e2 = __fmaf_rn(r2, inv, -tz);
    const float cost = dot3_ref(e0, e0, e1, e1, e2, e2);
    // 2x2 LM system (:202-215)
    const float A00 = __fadd_rn(dot3_ref(gx0, gx0, gx1, gx1, gx2, gx2), s.lambda);
    const float A01 = dot3_ref(gx0, gy0, gx1, gy1, gx2, gy2);
    const float A11 = __fadd_rn(dot3_ref(gy0, gy0, gy1, gy1, gy2, gy2), s.lambda);
    const float s0 = dot3_ref(gx0, e0, gx1, e1, gx2, e2);  // = -b0
    const float s1 = dot3_ref(gy0, e0, gy1, e1, gy2, e2);  // = -b1
    const float det_inv = __frcp_rn(__fmaf_rn(A00, A11, -__fmul_rn(A01, A01)));
    const float xu = __fmaf_rn(A01, s1, -__fmul_rn(s0, A11));  // A11*b0 - A01*b1
    const float xv = __fmaf_rn(A01, s0, -__fmul_rn(s1, A00));  // -A01*b0 + A00*b1
    const float u_new = fminf(fmaxf(__fmaf_rn(xu, det_inv, s.u), 1.0f), (float)(w - 2));
    const float v_new = fminf(fmaxf(__fmaf_rn(xv, det_inv, s.v), 1.0f), (float)(h - 2));

    // candidate cost (:225-256)
    const Bil Bn = bil_setup(u_new, v_new, w);
    float4 a11, a12, a21, a22;
    S.ray(Bn, a11, a12, a21, a22);
    const float t0 = bil(Bn, a11.x, a12.x, a21.x, a22.x);
    const float t1 = bil(Bn, a11.y, a12.y, a21.y, a22.y);
    const float t2 = bil(Bn, a11.z, a12.z, a21.z, a22.z);
    inv = __frcp_rn(__fsqrt_rn(dot3_ref(t0, t0, t1, t1, t2, t2)));
    const float f0 = __fmaf_rn(t0, inv, -tx), f1 = __fmaf_rn(t1, inv, -ty), f2 = __fmaf_rn(t2, inv, -tz);
    const float new_cost = dot3_ref(f0, f0, f1, f1, f2, f2);
    if (new_cost < cost) {  // accept (:259-264): the ray sampled at the candidate becomes the cached ray
      s.u = u_new;
      s.v = v_new;
      s.lambda = __double2float_rn((double)s.lambda * 0.1);
      s.conv = new_cost < cost_thresh;
      s.r0 = t0; s.r1 = t1; s.r2 = t2;
      s.gx0 = bil(Bn, a11.w, a12.w, a21.w, a22.w);
      s.B = Bn;
      s.have_r = true;
      s.have_g = false;
    } else {  // reject (:265-268): same (u,v) next time, everything cached
      s.lambda = __fmul_rn(s.lambda, 10.0f);
      s.conv = cost < cost_thresh;
      s.have_r = true;
      s.have_g = true;
    }
  }
}

__device__ __forceinline__ void iter_proj_core(const float4* __restrict__ p0, const float4* __restrict__ p1,
                                               const float* __restrict__ p2, float tx, float ty, float tz, int w, int h,
                                               int max_iter, float lambda_init, float cost_thresh, float& u, float& v,
                                               bool& conv_out) {
  IterProjState s;
  iter_proj_begin(s, u, v, w, lambda_init);
  const GlobalRaySampler S{p0, p1, p2, w};
  iter_proj_steps(s, S, tx, ty, tz, w, h, 0, max_iter, cost_thresh);
  u = s.u;
  v = s.v;
  conv_out = s.conv;
}

// One query of refine_matches in fp16 (matching_kernels.cu:25-81): query descriptor q (fp16 pairs), descriptor image of
// the query's batch item, start (u0, v0); returns the refined pixel.  Rounding after every multiply and add (see
// matching.cu).  The image is addressed as img[pixel * pstride + chunk * cstride] in 16-byte units: (NV, 1) for the
// reference's (h,w,fdim) layout.
template <int FDIM>
__device__ __forceinline__ void load_query_f16(const uint4* __restrict__ D21, long long gid, __half2 (&q)[FDIM / 2]) {
  constexpr int NV = FDIM / 8;
#pragma unroll
  for (int i = 0; i < NV; i++) {
    const uint4 t = __ldg(D21 + gid * NV + i);
    q[4 * i + 0] = *reinterpret_cast<const __half2*>(&t.x);
    q[4 * i + 1] = *reinterpret_cast<const __half2*>(&t.y);
    q[4 * i + 2] = *reinterpret_cast<const __half2*>(&t.z);
    q[4 * i + 3] = *reinterpret_cast<const __half2*>(&t.w);
  }
}

template <int FDIM>
__device__ __forceinline__ void load_query_f32(const float* __restrict__ D21, long long gid, __half2 (&q)[FDIM / 2]) {
#pragma unroll
  for (int i = 0; i < FDIM / 4; i++) {  // .half() of the reference glue (matching.py:81): round to nearest even
    const float4 t = __ldg(reinterpret_cast<const float4*>(D21 + gid * FDIM) + i);
    q[2 * i + 0] = __floats2half2_rn(t.x, t.y);
    q[2 * i + 1] = __floats2half2_rn(t.z, t.w);
  }
}

// fp16 score of one candidate: a product and a sum rounded after every operation, in descriptor order
template <int NV>
__device__ __forceinline__ float score_f16(const __half2* q, const uint4* t) {
  __half score = __ushort_as_half(0);
#pragma unroll
  for (int k = 0; k < NV; k++) {
    const __half2 a0 = *reinterpret_cast<const __half2*>(&t[k].x);
    const __half2 a1 = *reinterpret_cast<const __half2*>(&t[k].y);
    const __half2 a2 = *reinterpret_cast<const __half2*>(&t[k].z);
    const __half2 a3 = *reinterpret_cast<const __half2*>(&t[k].w);
    __half2 pr;
    pr = __hmul2_rn(q[4 * k + 0], a0);
    score = __hadd_rn(score, __low2half(pr));
    score = __hadd_rn(score, __high2half(pr));
    pr = __hmul2_rn(q[4 * k + 1], a1);
    score = __hadd_rn(score, __low2half(pr));
    score = __hadd_rn(score, __high2half(pr));
    pr = __hmul2_rn(q[4 * k + 2], a2);
    score = __hadd_rn(score, __low2half(pr));
    score = __hadd_rn(score, __high2half(pr));
    pr = __hmul2_rn(q[4 * k + 3], a3);
    score = __hadd_rn(score, __low2half(pr));
    score = __hadd_rn(score, __high2half(pr));
  }
  return __half2float(score);
}

// NJ candidates at once, their rounding chains interleaved: each candidate's score is the same serial sequence of
// 2 * FDIM/2 rounded additions as in score_f16 (so the results are bit-identical), but the NJ chains are independent and
// are issued round-robin -- one chain alone is 24-48 dependent HADD2 with nothing to overlap them.
template <int NV, int NJ>
__device__ __forceinline__ void score_f16_multi(const __half2* q, const uint4 (&t)[NJ][NV], float (&out)[NJ]) {
  __half sc[NJ];
#pragma unroll
  for (int j = 0; j < NJ; j++) sc[j] = __ushort_as_half(0);
#pragma unroll
  for (int k = 0; k < NV; k++) {
#pragma unroll
    for (int c = 0; c < 4; c++) {
      __half2 pr[NJ];
#pragma unroll
      for (int j = 0; j < NJ; j++) {
        const unsigned wv = c == 0 ? t[j][k].x : c == 1 ? t[j][k].y : c == 2 ? t[j][k].z : t[j][k].w;
        pr[j] = __hmul2_rn(q[4 * k + c], *reinterpret_cast<const __half2*>(&wv));
      }
#pragma unroll
      for (int j = 0; j < NJ; j++) sc[j] = __hadd_rn(sc[j], __low2half(pr[j]));
#pragma unroll
      for (int j = 0; j < NJ; j++) sc[j] = __hadd_rn(sc[j], __high2half(pr[j]));
    }
  }
#pragma unroll
  for (int j = 0; j < NJ; j++) out[j] = __half2float(sc[j]);
}

// RADIUS > 0: compile-time radius; the 2*RADIUS+1 candidates of one scan column are fetched together (all their loads
// in flight before the first score is reduced) -- the kernel is bound by dependent L1/L2 round trips, one candidate
// at a time, otherwise.  RADIUS == 0: run-time radius, one candidate at a time.  Scan order and the strict-greater
// argmax are the reference's (u offset outer, v offset inner, matching_kernels.cu:51-69).
template <int FDIM, int RADIUS>
__device__ __forceinline__ void refine_core_f16(const __half2 (&q)[FDIM / 2], const uint4* __restrict__ img, int pstride,
                                                int cstride, int h, int w, int radius, int dilation_max, int u0, int v0,
                                                int& u_out, int& v_out) {
  constexpr int NV = FDIM / 8;  // uint4 per descriptor
  int u_new = u0, v_new = v0;
  float max_score = 0.0f;  // cuda::std::numeric_limits<c10::Half>::min() == Half() == +0 (matching_kernels.cu:47)
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    const int diam = 2 * rd + 1;
    for (int i = 0; i < diam; i += d) {
      const int u = u0 - rd + i;
      if (u < 0 || u >= w) continue;
      if (RADIUS > 0) {
        constexpr int NJ = 2 * RADIUS + 1;
        uint4 t[NJ][NV];
        bool in[NJ];
#pragma unroll
        for (int jj = 0; jj < NJ; jj++) {
          const int v = v0 - rd + jj * d;
          in[jj] = v >= 0 && v < h;
          const uint4* c = img + (size_t)((in[jj] ? v : 0) * w + u) * pstride;
#pragma unroll
          for (int k = 0; k < NV; k++) t[jj][k] = in[jj] ? __ldg(c + (size_t)k * cstride) : make_uint4(0, 0, 0, 0);
        }
        float sc[NJ];
        score_f16_multi<NV, NJ>(q, t, sc);
#pragma unroll
        for (int jj = 0; jj < NJ; jj++) {
          if (in[jj] && sc[jj] > max_score) {
            max_score = sc[jj];
            u_new = u;
            v_new = v0 - rd + jj * d;
          }
        }
      } else {
        for (int j = 0; j < diam; j += d) {
          const int v = v0 - rd + j;
          if (v < 0 || v >= h) continue;
          const uint4* c = img + (size_t)(v * w + u) * pstride;
          uint4 t[NV];
#pragma unroll
          for (int k = 0; k < NV; k++) t[k] = __ldg(c + (size_t)k * cstride);
          const float s = score_f16<NV>(q, t);
          if (s > max_score) {
            max_score = s;
            u_new = u;
            v_new = v;
          }
        }
      }
    }
    u0 = u_new;
    v0 = v_new;
  }
  u_out = u_new;
  v_out = v_new;
}

}  // namespace m3r
