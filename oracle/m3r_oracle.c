/*
 * TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C) of the reference's dense backend
 * `mast3r_slam_backends` (yshhdec/mast3r-slam-ysh @ 74352e88).  It is the parity oracle for tests/,
 * __graft_entry__.smoke() and the cpu_baseline / --impl reference legs of bench.py.  The product
 * package (mast3r_slam_ysh_b200/) never imports, links or executes anything in this directory.
 *
 * Pinning status: the reference holds NO golden vectors / known-answer tests for this path
 * (SURVEY.md section 4, 8c).  This restatement is therefore pinned against outputs of the reference's own
 * CUDA kernels (oracle/_ref, compiled from /root/reference by oracle/build_ref.py and run on the
 * B200 box by tests/golden/make_golden.py); the vectors it produced are committed under
 * tests/golden/.  Until those vectors exist a checkout is "parity unpinned".
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/mast3r_slam/backend/src/).  Arithmetic mirrors the reference's mix of float and
 * double sub-expressions; it is compiled with -ffp-contract=off, so the only systematic difference
 * from the reference binary is nvcc's FMA contraction (covered by the stated tolerances).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------------
 * fp16 <-> fp32 (IEEE binary16, round-to-nearest-even), bit-level.  Mirrors c10::Half conversions
 * used by refine_matches_kernel<c10::Half> (matching_kernels.cu:60-63).
 * ---------------------------------------------------------------------------------------------- */
static inline float orc_h2f(uint16_t h) {
  uint32_t sign = ((uint32_t)h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1fu;
  uint32_t man = h & 0x3ffu;
  uint32_t bits;
  if (exp == 0) {
    if (man == 0) {
      bits = sign;
    } else { /* subnormal: normalise */
      int e = -1;
      do { man <<= 1; e++; } while (!(man & 0x400u));
      man &= 0x3ffu;
      bits = sign | ((uint32_t)(127 - 15 - e) << 23) | (man << 13);
    }
  } else if (exp == 31) {
    bits = sign | 0x7f800000u | (man << 13);
  } else {
    bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
  }
  float f;
  memcpy(&f, &bits, 4);
  return f;
}

static inline uint16_t orc_f2h(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  uint16_t sign = (uint16_t)((x >> 16) & 0x8000u);
  uint32_t ax = x & 0x7fffffffu;
  if (ax >= 0x7f800000u) { /* inf / nan */
    return (uint16_t)(sign | 0x7c00u | ((ax > 0x7f800000u) ? 0x200u : 0));
  }
  if (ax >= 0x477ff000u) { /* >= 65520 rounds to inf */
    return (uint16_t)(sign | 0x7c00u);
  }
  if (ax < 0x33000001u) { /* <= 2^-25 rounds to zero (ties-to-even at exactly 2^-25) */
    return sign;
  }
  int32_t e = (int32_t)(ax >> 23) - 127;
  uint32_t m = (ax & 0x7fffffu) | 0x800000u; /* 24-bit significand */
  uint32_t shift;
  uint32_t hexp;
  if (e < -14) { /* subnormal result */
    shift = (uint32_t)(13 + (-14 - e));
    hexp = 0;
  } else {
    shift = 13;
    hexp = (uint32_t)(e + 15);
  }
  uint32_t q = m >> shift;
  uint32_t rem = m & ((1u << shift) - 1u);
  uint32_t half = 1u << (shift - 1);
  if (rem > half || (rem == half && (q & 1u))) q++;
  uint32_t out;
  if (hexp == 0) {
    out = q; /* may carry into exponent 1: correct */
  } else {
    out = ((hexp - 1) << 10) + q; /* q includes the hidden bit (0x400) */
  }
  return (uint16_t)(sign | out);
}

ORC_API void orc_f32_to_f16(const float* src, uint16_t* dst, int64_t n) {
  for (int64_t i = 0; i < n; i++) dst[i] = orc_f2h(src[i]);
}
ORC_API void orc_f16_to_f32(const uint16_t* src, float* dst, int64_t n) {
  for (int64_t i = 0; i < n; i++) dst[i] = orc_h2f(src[i]);
}

/* ------------------------------------------------------------------------------------------------
 * refine_matches  (matching_kernels.cu:25-81 kernel, :84-116 launcher)
 *   D11 (b,h,w,fdim), D21 (b,n,fdim), p1 (b,n,2) int64 [u,v] -> p1_new (b,n,2)
 *   half: score accumulates as c10::Half: each multiply and each add is computed in fp32 and
 *   rounded to fp16 (:60-63); max_score starts at cuda::std::numeric_limits<c10::Half>::min() which is the
 *   unspecialised primary template -> Half() == +0 (:47); compare is strict '>' in float (:65).
 * ---------------------------------------------------------------------------------------------- */
ORC_API int orc_refine_matches_f16(const uint16_t* D11, const uint16_t* D21, const int64_t* p1, int64_t* p1_new,
                                   int b, int h, int w, int n, int fdim, int radius, int dilation_max) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int bi = 0; bi < b; bi++) {
    for (int q = 0; q < n; q++) {
      const int64_t* p = p1 + ((int64_t)bi * n + q) * 2;
      const uint16_t* d21 = D21 + ((int64_t)bi * n + q) * fdim;
      long u0 = p[0], v0 = p[1];
      float max_score = 0.0f; /* Half() */
      long u_new = u0, v_new = v0;
      for (int d = dilation_max; d > 0; d--) {
        const int rd = radius * d;
        const int diam = 2 * rd + 1;
        for (int i = 0; i < diam; i += d) {
          for (int j = 0; j < diam; j += d) {
            const long u = u0 - rd + i;
            const long v = v0 - rd + j;
            if (v >= 0 && v < h && u >= 0 && u < w) {
              const uint16_t* d11 = D11 + (((int64_t)bi * h + v) * w + u) * fdim;
              uint16_t score = 0;
              for (int k = 0; k < fdim; k++) {
                uint16_t prod = orc_f2h(orc_h2f(d21[k]) * orc_h2f(d11[k]));
                score = orc_f2h(orc_h2f(score) + orc_h2f(prod));
              }
              float s = orc_h2f(score);
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
      p1_new[((int64_t)bi * n + q) * 2 + 0] = u_new;
      p1_new[((int64_t)bi * n + q) * 2 + 1] = v_new;
    }
  }
  return 0;
}

/* float instantiation of the same template: scalar_t = float, max_score = FLT_MIN (:47). */
ORC_API int orc_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1, int64_t* p1_new,
                                   int b, int h, int w, int n, int fdim, int radius, int dilation_max) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int bi = 0; bi < b; bi++) {
    for (int q = 0; q < n; q++) {
      const int64_t* p = p1 + ((int64_t)bi * n + q) * 2;
      const float* d21 = D21 + ((int64_t)bi * n + q) * fdim;
      long u0 = p[0], v0 = p[1];
      float max_score = FLT_MIN;
      long u_new = u0, v_new = v0;
      for (int d = dilation_max; d > 0; d--) {
        const int rd = radius * d;
        const int diam = 2 * rd + 1;
        for (int i = 0; i < diam; i += d) {
          for (int j = 0; j < diam; j += d) {
            const long u = u0 - rd + i;
            const long v = v0 - rd + j;
            if (v >= 0 && v < h && u >= 0 && u < w) {
              const float* d11 = D11 + (((int64_t)bi * h + v) * w + u) * fdim;
              float score = 0.0f;
              for (int k = 0; k < fdim; k++) score += d21[k] * d11[k];
              if (score > max_score) {
                max_score = score;
                u_new = u;
                v_new = v;
              }
            }
          }
        }
        u0 = u_new;
        v0 = v_new;
      }
      p1_new[((int64_t)bi * n + q) * 2 + 0] = u_new;
      p1_new[((int64_t)bi * n + q) * 2 + 1] = v_new;
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * iter_proj  (matching_kernels.cu:119-275 kernel, :279-316 launcher)
 *   rays (b,h,w,9) = [ray, d ray/du, d ray/dv]; pts (b,n,3) unit; p_init (b,n,2) float [u,v]
 *   -> p_new (b,n,2), converged (b,n) bool.  Double sub-expressions mirrored: (1.0-du), 1.0/r_norm,
 *   1.0/det, lambda*=0.1 / 10.0 (:162-164,187,213,262,266).
 * ---------------------------------------------------------------------------------------------- */
static inline float orc_clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

ORC_API int orc_iter_proj(const float* rays, const float* pts, const float* p_init, float* p_new, uint8_t* converged,
                          int b, int h, int w, int n, int max_iter, float lambda_init, float cost_thresh) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int bi = 0; bi < b; bi++) {
    for (int q = 0; q < n; q++) {
      const float* img = rays + (int64_t)bi * h * w * 9;
      const float* tgt = pts + ((int64_t)bi * n + q) * 3;
      float u = p_init[((int64_t)bi * n + q) * 2 + 0];
      float v = p_init[((int64_t)bi * n + q) * 2 + 1];
      u = orc_clampf(u, 1, (float)(w - 2));
      v = orc_clampf(v, 1, (float)(h - 2));
      float r[3], gx[3], gy[3], err[3];
      float lambda = lambda_init;
      uint8_t conv = 0; /* torch::zeros (:300) */
      for (int it = 0; it < max_iter; it++) {
        int u11 = (int)floorf(u);
        int v11 = (int)floorf(v);
        float du = u - (float)u11;
        float dv = v - (float)v11;
        float w11 = du * dv;
        float w12 = (float)((1.0 - du) * dv);
        float w21 = (float)(du * (1.0 - dv));
        float w22 = (float)((1.0 - du) * (1.0 - dv));
        const float* r11 = img + ((int64_t)(v11 + 1) * w + (u11 + 1)) * 9;
        const float* r12 = img + ((int64_t)(v11 + 1) * w + u11) * 9;
        const float* r21 = img + ((int64_t)v11 * w + (u11 + 1)) * 9;
        const float* r22 = img + ((int64_t)v11 * w + u11) * 9;
        for (int j = 0; j < 3; j++) r[j] = w11 * r11[j] + w12 * r12[j] + w21 * r21[j] + w22 * r22[j];
        for (int j = 3; j < 6; j++) gx[j - 3] = w11 * r11[j] + w12 * r12[j] + w21 * r21[j] + w22 * r22[j];
        for (int j = 6; j < 9; j++) gy[j - 6] = w11 * r11[j] + w12 * r12[j] + w21 * r21[j] + w22 * r22[j];
        float r_norm = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
        float r_norm_inv = (float)(1.0 / r_norm);
        for (int j = 0; j < 3; j++) r[j] *= r_norm_inv;
        for (int j = 0; j < 3; j++) err[j] = r[j] - tgt[j];
        float cost = err[0] * err[0] + err[1] * err[1] + err[2] * err[2];
        float A00 = gx[0] * gx[0] + gx[1] * gx[1] + gx[2] * gx[2];
        float A01 = gx[0] * gy[0] + gx[1] * gy[1] + gx[2] * gy[2];
        float A11 = gy[0] * gy[0] + gy[1] * gy[1] + gy[2] * gy[2];
        float b0 = -(err[0] * gx[0] + err[1] * gx[1] + err[2] * gx[2]);
        float b1 = -(err[0] * gy[0] + err[1] * gy[1] + err[2] * gy[2]);
        A00 += lambda;
        A11 += lambda;
        float det_inv = (float)(1.0 / (A00 * A11 - A01 * A01));
        float delta_u = det_inv * (A11 * b0 - A01 * b1);
        float delta_v = det_inv * (-A01 * b0 + A00 * b1);
        float u_new = u + delta_u;
        float v_new = v + delta_v;
        u_new = orc_clampf(u_new, 1, (float)(w - 2));
        v_new = orc_clampf(v_new, 1, (float)(h - 2));
        u11 = (int)floorf(u_new);
        v11 = (int)floorf(v_new);
        du = u_new - u11;
        dv = v_new - v11;
        w11 = du * dv;
        w12 = (float)((1.0 - du) * dv);
        w21 = (float)(du * (1.0 - dv));
        w22 = (float)((1.0 - du) * (1.0 - dv));
        r11 = img + ((int64_t)(v11 + 1) * w + (u11 + 1)) * 9;
        r12 = img + ((int64_t)(v11 + 1) * w + u11) * 9;
        r21 = img + ((int64_t)v11 * w + (u11 + 1)) * 9;
        r22 = img + ((int64_t)v11 * w + u11) * 9;
        for (int j = 0; j < 3; j++) r[j] = w11 * r11[j] + w12 * r12[j] + w21 * r21[j] + w22 * r22[j];
        r_norm = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
        r_norm_inv = (float)(1.0 / r_norm);
        for (int j = 0; j < 3; j++) r[j] *= r_norm_inv;
        for (int j = 0; j < 3; j++) err[j] = r[j] - tgt[j];
        float new_cost = err[0] * err[0] + err[1] * err[1] + err[2] * err[2];
        if (new_cost < cost) {
          u = u_new;
          v = v_new;
          lambda = (float)(lambda * 0.1);
          conv = new_cost < cost_thresh;
        } else {
          lambda = (float)(lambda * 10.0);
          conv = cost < cost_thresh;
        }
      }
      p_new[((int64_t)bi * n + q) * 2 + 0] = u;
      p_new[((int64_t)bi * n + q) * 2 + 1] = v;
      converged[(int64_t)bi * n + q] = conv;
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * Sim(3) helpers (gn_kernels.cu:172-413)
 * ---------------------------------------------------------------------------------------------- */
#define ORC_EPS 1e-6 /* gn_kernels.cu:34 (double literal) */

static inline float orc_huber(float r) { /* :172-175, k = 1.345 compared/divided in double */
  const float r_abs = fabsf(r);
  return (float)(r_abs < 1.345 ? 1.0 : 1.345 / r_abs);
}

static void orc_quat_comp(const float* qi, const float* qj, float* out) { /* :178-184 */
  out[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  out[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  out[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  out[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}

static void orc_actSO3(const float* q, const float* X, float* Y) { /* :195-205 (Y may alias X) */
  float uv[3];
  uv[0] = (float)(2.0 * (q[1] * X[2] - q[2] * X[1]));
  uv[1] = (float)(2.0 * (q[2] * X[0] - q[0] * X[2]));
  uv[2] = (float)(2.0 * (q[0] * X[1] - q[1] * X[0]));
  float y0 = X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  float y1 = X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  float y2 = X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
  Y[0] = y0; Y[1] = y1; Y[2] = y2;
}

static void orc_actSim3(const float* t, const float* q, const float* s, const float* X, float* Y) { /* :207-219 */
  orc_actSO3(q, X, Y);
  Y[0] *= s[0]; Y[1] *= s[0]; Y[2] *= s[0];
  Y[0] += t[0]; Y[1] += t[1]; Y[2] += t[2];
}

static void orc_relSim3(const float* ti, const float* qi, const float* si, const float* tj, const float* qj,
                        const float* sj, float* tij, float* qij, float* sij) { /* :252-272 */
  float si_inv = (float)(1.0 / si[0]);
  sij[0] = si_inv * sj[0];
  float qi_inv[4] = {-qi[0], -qi[1], -qi[2], qi[3]};
  orc_quat_comp(qi_inv, qj, qij);
  tij[0] = tj[0] - ti[0];
  tij[1] = tj[1] - ti[1];
  tij[2] = tj[2] - ti[2];
  /* NOTE: the reference calls actSO3(qi_inv, tij, tij) IN PLACE (:270): Y[0] is written before
   * Y[1], Y[2] read X[0].  uv[] is computed from the original X first, but the Y[] lines re-read
   * X[0..2] one by one; X[0] is only read by Y[0], X[1] by Y[1], X[2] by Y[2], so aliasing is
   * benign and the result equals the out-of-place value. */
  orc_actSO3(qi_inv, tij, tij);
  tij[0] *= si_inv; tij[1] *= si_inv; tij[2] *= si_inv;
}

static void orc_apply_Sim3_adj_inv(const float* t, const float* q, const float* s, const float* X, float* Y) { /* :277-297 */
  const float s_inv = (float)(1.0 / s[0]);
  float Ra[3];
  orc_actSO3(q, &X[0], Ra);
  Y[0] = s_inv * Ra[0];
  Y[1] = s_inv * Ra[1];
  Y[2] = s_inv * Ra[2];
  orc_actSO3(q, &X[3], &Y[3]);
  Y[3] += s_inv * (t[1] * Ra[2] - t[2] * Ra[1]);
  Y[4] += s_inv * (t[2] * Ra[0] - t[0] * Ra[2]);
  Y[5] += s_inv * (t[0] * Ra[1] - t[1] * Ra[0]);
  Y[6] = X[6] + (s_inv * (t[0] * Ra[0] + t[1] * Ra[1] + t[2] * Ra[2]));
}

static void orc_expSO3(const float* phi, float* q) { /* :299-321 */
  float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float imag, real;
  if (theta_sq < ORC_EPS) {
    float theta_p4 = theta_sq * theta_sq;
    imag = (float)(0.5 - (1.0 / 48.0) * theta_sq + (1.0 / 3840.0) * theta_p4);
    real = (float)(1.0 - (1.0 / 8.0) * theta_sq + (1.0 / 384.0) * theta_p4);
  } else {
    float theta = sqrtf(theta_sq);
    imag = sinf((float)(0.5 * theta)) / theta;
    real = cosf((float)(0.5 * theta));
  }
  q[0] = imag * phi[0];
  q[1] = imag * phi[1];
  q[2] = imag * phi[2];
  q[3] = real;
}

static void orc_cross_inplace(const float* a, float* b) { /* :229-240 */
  float x0 = a[1] * b[2] - a[2] * b[1];
  float x1 = a[2] * b[0] - a[0] * b[2];
  float x2 = a[0] * b[1] - a[1] * b[0];
  b[0] = x0; b[1] = x1; b[2] = x2;
}

static void orc_expSim3(const float* xi, float* t, float* q, float* s) { /* :323-390 */
  float tau[3] = {xi[0], xi[1], xi[2]};
  float phi[3] = {xi[3], xi[4], xi[5]};
  float sigma = xi[6];
  float scale = expf(sigma);
  orc_expSO3(phi, q);
  s[0] = scale;
  float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float theta = sqrtf(theta_sq);
  float A, B, C;
  const float one = 1.0f, half = 0.5f;
  if (fabsf(sigma) < ORC_EPS) {
    C = one;
    if (fabsf(theta) < ORC_EPS) {
      A = half;
      B = (float)(1.0 / 6.0);
    } else {
      A = (one - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - one) / sigma;
    if (fabsf(theta) < ORC_EPS) {
      float sigma_sq = sigma * sigma;
      A = ((sigma - one) * scale + one) / sigma_sq;
      B = (scale * half * sigma_sq + scale - one - sigma * scale) / (sigma_sq * sigma);
    } else {
      float a = scale * sinf(theta);
      float b = scale * cosf(theta);
      float c = theta_sq + sigma * sigma;
      A = (a * sigma + (one - b) * theta) / (theta * c);
      B = (C - ((b - one) * sigma + a * theta) / (c)) / (theta_sq);
    }
  }
  t[0] = C * tau[0];
  t[1] = C * tau[1];
  t[2] = C * tau[2];
  orc_cross_inplace(phi, tau);
  t[0] += A * tau[0];
  t[1] += A * tau[1];
  t[2] += A * tau[2];
  orc_cross_inplace(phi, tau);
  t[0] += B * tau[0];
  t[1] += B * tau[1];
  t[2] += B * tau[2];
}

static void orc_retrSim3(const float* xi, const float* t, const float* q, const float* s, float* t1, float* q1, float* s1) { /* :392-413 */
  float dt[3] = {0, 0, 0}, dq[4] = {0, 0, 0, 1}, ds[1] = {0};
  orc_expSim3(xi, dt, dq, ds);
  orc_quat_comp(dq, q, q1);
  orc_actSO3(dq, t, t1);
  t1[0] *= ds[0]; t1[1] *= ds[0]; t1[2] *= ds[0];
  t1[0] += dt[0]; t1[1] += dt[1]; t1[2] += dt[2];
  s1[0] = ds[0] * s[0];
}

/* pose_retr_kernel (:415-453): poses (num_poses,8) in place, dx (num_poses-num_fix,7) */
ORC_API void orc_pose_retr(float* poses, const float* dx, int num_poses, int num_fix) {
  for (int k = num_fix; k < num_poses; k++) {
    float* p = poses + (int64_t)k * 8;
    float t[3] = {p[0], p[1], p[2]}, q[4] = {p[3], p[4], p[5], p[6]}, s[1] = {p[7]};
    float t1[3], q1[4], s1[1];
    orc_retrSim3(dx + (int64_t)(k - num_fix) * 7, t, q, s, t1, q1, s1);
    p[0] = t1[0]; p[1] = t1[1]; p[2] = t1[2];
    p[3] = q1[0]; p[4] = q1[1]; p[5] = q1[2]; p[6] = q1[3];
    p[7] = s1[0];
  }
}

/* exposed for unit tests of the group helpers */
ORC_API void orc_rel_sim3(const float* Ti, const float* Tj, float* Tij) {
  orc_relSim3(Ti, Ti + 3, Ti + 7, Tj, Tj + 3, Tj + 7, Tij, Tij + 3, Tij + 7);
}
ORC_API void orc_act_sim3(const float* T, const float* X, float* Y, int n) {
  for (int i = 0; i < n; i++) orc_actSim3(T, T + 3, T + 7, X + 3 * i, Y + 3 * i);
}
ORC_API void orc_adj_inv_row(const float* T, const float* X, float* Y) { orc_apply_Sim3_adj_inv(T, T + 3, T + 7, X, Y); }

/* ------------------------------------------------------------------------------------------------
 * Per-edge residual / Jacobian / Hessian blocks.
 *   kind 0: point_align_kernel (:455-723)   params = {sigma_point}
 *   kind 1: ray_align_kernel   (:813-1138)  params = {sigma_ray, sigma_dist}
 *   kind 2: calib_proj_kernel  (:1231-1543) params = {sigma_pixel, sigma_depth, fx, fy, cx, cy,
 *                                                     height, width, pixel_border, z_eps}
 *   robust 0 = Huber (reference); 1 = Tukey (nonlinear_optimizer.py:36-42; opt-in, not used by the reference kernels)
 *   accum 0: each edge's 105+14 sums accumulated in fp64 over pixels in index order (bounds the oracle's own error)
 *   accum 1: fp32, the reference's order: 256 strided per-thread partial sums (:31-32) followed by the
 *            shared-memory tree of blockReduce (:36-55) -- like-for-like with the reference binary.
 *   Outputs: Hs (4,E,7,7) = [H_ii,H_ij,H_ji,H_jj], gs (2,E,7) = [g_i,g_j]   (:1097-1137)
 * ---------------------------------------------------------------------------------------------- */
#define ORC_THREADS 256

static inline float orc_tukey(float r) {
  const float t = 4.6851f;
  float r_abs = fabsf(r);
  float tmp = 1.0f - (r_abs / t) * (r_abs / t);
  return r_abs < t ? tmp * tmp : 0.0f;
}
static inline float orc_robust(float r, int robust) { return robust ? orc_tukey(r) : orc_huber(r); }

/* the IRLS weight functions alone, for pinning against nonlinear_optimizer.py:28-42 (tests/golden/robust.npz) */
ORC_API int orc_robust_weights(const float* r, int64_t n, int robust, float* w) {
  for (int64_t i = 0; i < n; i++) w[i] = orc_robust(r[i], robust);
  return 0;
}

typedef struct {
  int nres;
  float err[4];
  float w[4];
  float J[4][7]; /* local rows ("Ji" before the adjoint is applied) */
} orc_rows_t;

static void orc_pixel_rows(int kind, const float* params, int robust, const float* tij, const float* qij,
                           const float* sij, const float* Xi, const float* Xj, int64_t ind_Xi, int valid_match_ind,
                           float q, float ci, float cj, float C_thresh, float Q_thresh, orc_rows_t* o) {
  float P[3];
  orc_actSim3(tij, qij, sij, Xj, P);
  memset(o->J, 0, sizeof(o->J));
  if (kind == 0) { /* points */
    const float sigma_point_inv = (float)(1.0 / params[0]);
    o->nres = 3;
    o->err[0] = P[0] - Xi[0];
    o->err[1] = P[1] - Xi[1];
    o->err[2] = P[2] - Xi[2];
    const int valid = valid_match_ind & (q > Q_thresh) & (ci > C_thresh) & (cj > C_thresh);
    const float sw = valid ? sigma_point_inv * sqrtf(q) : 0;
    const float wc = sw * sw;
    for (int k = 0; k < 3; k++) o->w[k] = orc_robust(sw * o->err[k], robust) * wc;
    float* J = o->J[0];
    J[0] = 1; J[4] = P[2]; J[5] = -P[1]; J[6] = P[0];
    J = o->J[1];
    J[1] = 1; J[3] = -P[2]; J[5] = P[0]; J[6] = P[1];
    J = o->J[2];
    J[2] = 1; J[3] = P[1]; J[4] = -P[0]; J[6] = P[2];
  } else if (kind == 1) { /* rays + distance */
    const float sigma_ray_inv = (float)(1.0 / params[0]);
    const float sigma_dist_inv = (float)(1.0 / params[1]);
    o->nres = 4;
    const float norm2_i = Xi[0] * Xi[0] + Xi[1] * Xi[1] + Xi[2] * Xi[2];
    const float norm1_i = sqrtf(norm2_i);
    const float norm1_i_inv = (float)(1.0 / norm1_i);
    float ri[3];
    for (int i = 0; i < 3; i++) ri[i] = norm1_i_inv * Xi[i];
    const float norm2_j = P[0] * P[0] + P[1] * P[1] + P[2] * P[2];
    const float norm1_j = sqrtf(norm2_j);
    const float norm1_j_inv = (float)(1.0 / norm1_j);
    float rj[3];
    for (int i = 0; i < 3; i++) rj[i] = norm1_j_inv * P[i];
    o->err[0] = rj[0] - ri[0];
    o->err[1] = rj[1] - ri[1];
    o->err[2] = rj[2] - ri[2];
    o->err[3] = norm1_j - norm1_i;
    const int valid = valid_match_ind & (q > Q_thresh) & (ci > C_thresh) & (cj > C_thresh);
    const float sw_ray = valid ? sigma_ray_inv * sqrtf(q) : 0;
    const float sw_dist = valid ? sigma_dist_inv * sqrtf(q) : 0;
    o->w[0] = orc_robust(sw_ray * o->err[0], robust);
    o->w[1] = orc_robust(sw_ray * o->err[1], robust);
    o->w[2] = orc_robust(sw_ray * o->err[2], robust);
    o->w[3] = orc_robust(sw_dist * o->err[3], robust);
    const float wc_ray = sw_ray * sw_ray, wc_dist = sw_dist * sw_dist;
    o->w[0] *= wc_ray; o->w[1] *= wc_ray; o->w[2] *= wc_ray; o->w[3] *= wc_dist;
    const float norm3_j_inv = norm1_j_inv / norm2_j;
    const float drx_dPx = norm1_j_inv - P[0] * P[0] * norm3_j_inv;
    const float dry_dPy = norm1_j_inv - P[1] * P[1] * norm3_j_inv;
    const float drz_dPz = norm1_j_inv - P[2] * P[2] * norm3_j_inv;
    const float drx_dPy = -P[0] * P[1] * norm3_j_inv;
    const float drx_dPz = -P[0] * P[2] * norm3_j_inv;
    const float dry_dPz = -P[1] * P[2] * norm3_j_inv;
    float* J = o->J[0];
    J[0] = drx_dPx; J[1] = drx_dPy; J[2] = drx_dPz; J[4] = rj[2]; J[5] = -rj[1];
    J = o->J[1];
    J[0] = drx_dPy; J[1] = dry_dPy; J[2] = dry_dPz; J[3] = -rj[2]; J[5] = rj[0];
    J = o->J[2];
    J[0] = drx_dPz; J[1] = dry_dPz; J[2] = drz_dPz; J[3] = rj[1]; J[4] = -rj[0];
    J = o->J[3];
    J[0] = rj[0]; J[1] = rj[1]; J[2] = rj[2]; J[6] = norm1_j;
  } else { /* calibrated projection + log depth */
    const float sigma_pixel_inv = (float)(1.0 / params[0]);
    const float sigma_depth_inv = (float)(1.0 / params[1]);
    const float fx = params[2], fy = params[3], cx = params[4], cy = params[5];
    const int height = (int)params[6], width = (int)params[7], pixel_border = (int)params[8];
    const float z_eps = params[9];
    o->nres = 3;
    const int u_target = (int)(ind_Xi % width);
    const int v_target = (int)(ind_Xi / width);
    const int valid_z = (P[2] > z_eps) && (Xi[2] > z_eps);
    const float zj_inv = valid_z ? (float)(1.0 / P[2]) : 0.0f;
    const float zj_log = valid_z ? logf(P[2]) : 0.0f;
    const float zi_log = valid_z ? logf(Xi[2]) : 0.0f;
    const float x_div_z = P[0] * zj_inv;
    const float y_div_z = P[1] * zj_inv;
    const float u = fx * x_div_z + cx;
    const float v = fy * y_div_z + cy;
    const int valid_u = (u > pixel_border) && (u < width - 1 - pixel_border);
    const int valid_v = (v > pixel_border) && (v < height - 1 - pixel_border);
    o->err[0] = u - u_target;
    o->err[1] = v - v_target;
    o->err[2] = zj_log - zi_log;
    const int valid = valid_match_ind & (q > Q_thresh) & (ci > C_thresh) & (cj > C_thresh) & valid_u & valid_v & valid_z;
    const float sw_pixel = valid ? sigma_pixel_inv * sqrtf(q) : 0;
    const float sw_depth = valid ? sigma_depth_inv * sqrtf(q) : 0;
    o->w[0] = orc_robust(sw_pixel * o->err[0], robust);
    o->w[1] = orc_robust(sw_pixel * o->err[1], robust);
    o->w[2] = orc_robust(sw_depth * o->err[2], robust);
    const float wc_pixel = sw_pixel * sw_pixel, wc_depth = sw_depth * sw_depth;
    o->w[0] *= wc_pixel; o->w[1] *= wc_pixel; o->w[2] *= wc_depth;
    float* J = o->J[0];
    J[0] = fx * zj_inv;
    J[2] = -fx * x_div_z * zj_inv;
    J[3] = -fx * x_div_z * y_div_z;
    J[4] = fx * (1 + x_div_z * x_div_z);
    J[5] = -fx * y_div_z;
    J = o->J[1];
    J[1] = fy * zj_inv;
    J[2] = -fy * y_div_z * zj_inv;
    J[3] = -fy * (1 + y_div_z * y_div_z);
    J[4] = fy * x_div_z * y_div_z;
    J[5] = fy * x_div_z;
    J = o->J[2];
    J[2] = zj_inv;
    J[3] = y_div_z;
    J[4] = -x_div_z;
    J[6] = 1.0f;
  }
}

static void orc_edge_block(int kind, const float* params, int robust, int accum, const float* Twc, const float* Xs,
                           const float* Cs, int ix, int jx, const int64_t* idx, const uint8_t* valid_match,
                           const float* Q, int npts, float C_thresh, float Q_thresh, double* hsum /*105*/,
                           double* visum /*7*/, double* vjsum /*7*/) {
  const float* Ti = Twc + (int64_t)ix * 8;
  const float* Tj = Twc + (int64_t)jx * 8;
  float tij[3], qij[4], sij[1];
  orc_relSim3(Ti, Ti + 3, Ti + 7, Tj, Tj + 3, Tj + 7, tij, qij, sij);
  const float* Xi_base = Xs + (int64_t)ix * npts * 3;
  const float* Xj_base = Xs + (int64_t)jx * npts * 3;
  const float* Ci_base = Cs + (int64_t)ix * npts;
  const float* Cj_base = Cs + (int64_t)jx * npts;
  const int nthr = accum ? ORC_THREADS : 1;
  float* hf = NULL; float* vif = NULL; float* vjf = NULL;
  double hd[105], vid[7], vjd[7];
  if (accum) {
    hf = (float*)calloc((size_t)ORC_THREADS * 105, sizeof(float));
    vif = (float*)calloc((size_t)ORC_THREADS * 7, sizeof(float));
    vjf = (float*)calloc((size_t)ORC_THREADS * 7, sizeof(float));
  } else {
    memset(hd, 0, sizeof(hd)); memset(vid, 0, sizeof(vid)); memset(vjd, 0, sizeof(vjd));
  }
  orc_rows_t rows;
  for (int64_t k = 0; k < npts; k++) {
    const int vm = valid_match[k] != 0;
    const int64_t ind_Xi = vm ? idx[k] : 0;
    const float* Xi = Xi_base + ind_Xi * 3;
    const float* Xj = Xj_base + k * 3;
    orc_pixel_rows(kind, params, robust, tij, qij, sij, Xi, Xj, ind_Xi, vm, Q[k], Ci_base[ind_Xi], Cj_base[k],
                   C_thresh, Q_thresh, &rows);
    const int t = (int)(k % nthr);
    for (int r = 0; r < rows.nres; r++) {
      float Jx[14];
      orc_apply_Sim3_adj_inv(Ti, Ti + 3, Ti + 7, rows.J[r], &Jx[7]);
      for (int n = 0; n < 7; n++) Jx[n] = -Jx[7 + n];
      const float w = rows.w[r], e = rows.err[r];
      int l = 0;
      if (accum) {
        float* h = hf + (size_t)t * 105;
        for (int n = 0; n < 14; n++)
          for (int m = 0; m <= n; m++) { h[l] += w * Jx[n] * Jx[m]; l++; }
        for (int n = 0; n < 7; n++) {
          vif[t * 7 + n] += w * e * Jx[n];
          vjf[t * 7 + n] += w * e * Jx[7 + n];
        }
      } else {
        for (int n = 0; n < 14; n++)
          for (int m = 0; m <= n; m++) { hd[l] += (double)(w * Jx[n] * Jx[m]); l++; }
        for (int n = 0; n < 7; n++) {
          vid[n] += (double)(w * e * Jx[n]);
          vjd[n] += (double)(w * e * Jx[7 + n]);
        }
      }
    }
  }
  if (accum) {
    /* blockReduce (:45-55) + warpReduce (:36-43): s[t] += s[t+128]; +64; +32; +16; ... +1, fp32 */
    float s[ORC_THREADS];
    for (int l = 0; l < 105 + 14; l++) {
      for (int t = 0; t < ORC_THREADS; t++)
        s[t] = l < 105 ? hf[(size_t)t * 105 + l] : (l < 112 ? vif[t * 7 + (l - 105)] : vjf[t * 7 + (l - 112)]);
      for (int stride = 128; stride >= 1; stride >>= 1)
        for (int t = 0; t < stride; t++) s[t] += s[t + stride];
      if (l < 105) hsum[l] = s[0]; else if (l < 112) visum[l - 105] = s[0]; else vjsum[l - 112] = s[0];
    }
    free(hf); free(vif); free(vjf);
  } else {
    memcpy(hsum, hd, sizeof(hd)); memcpy(visum, vid, sizeof(vid)); memcpy(vjsum, vjd, sizeof(vjd));
  }
}

ORC_API int orc_gn_edge_blocks(int kind, const float* params, int robust, int accum, const float* Twc,
                               const float* Xs, const float* Cs, const int64_t* ii_edge, const int64_t* jj_edge,
                               const int64_t* idx, const uint8_t* valid_match, const float* Q, int num_kf, int npts,
                               int E, float C_thresh, float Q_thresh, float* Hs /*4,E,7,7*/, float* gs /*2,E,7*/) {
  (void)num_kf;
#pragma omp parallel for schedule(dynamic, 1)
  for (int e = 0; e < E; e++) {
    double h[105], vi[7], vj[7];
    orc_edge_block(kind, params, robust, accum, Twc, Xs, Cs, (int)ii_edge[e], (int)jj_edge[e],
                   idx + (int64_t)e * npts, valid_match + (int64_t)e * npts, Q + (int64_t)e * npts, npts, C_thresh,
                   Q_thresh, h, vi, vj);
    float* H0 = Hs + ((int64_t)0 * E + e) * 49;
    float* H1 = Hs + ((int64_t)1 * E + e) * 49;
    float* H2 = Hs + ((int64_t)2 * E + e) * 49;
    float* H3 = Hs + ((int64_t)3 * E + e) * 49;
    for (int n = 0; n < 7; n++) {
      gs[((int64_t)0 * E + e) * 7 + n] = (float)vi[n];
      gs[((int64_t)1 * E + e) * 7 + n] = (float)vj[n];
    }
    int l = 0;
    for (int n = 0; n < 14; n++) {
      for (int m = 0; m <= n; m++) {
        const float v = (float)h[l];
        if (n < 7 && m < 7) {
          H0[n * 7 + m] = v; H0[m * 7 + n] = v;
        } else if (n >= 7 && m < 7) {
          H1[m * 7 + (n - 7)] = v; H2[(n - 7) * 7 + m] = v;
        } else {
          H3[(n - 7) * 7 + (m - 7)] = v; H3[(m - 7) * 7 + (n - 7)] = v;
        }
        l++;
      }
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * Index prep (gn_kernels.cu:161-170): sorted unique of cat(ii,jj); searchsorted(unique, x) - pin.
 * ---------------------------------------------------------------------------------------------- */
static int orc_cmp_i64(const void* a, const void* b) {
  int64_t x = *(const int64_t*)a, y = *(const int64_t*)b;
  return (x > y) - (x < y);
}

ORC_API int orc_index_prep(const int64_t* ii, const int64_t* jj, int E, int64_t* uniq /*2E*/, int64_t* ii_edge,
                           int64_t* jj_edge) {
  int64_t* all = (int64_t*)malloc(sizeof(int64_t) * 2 * (size_t)(E > 0 ? E : 1));
  memcpy(all, ii, sizeof(int64_t) * E);
  memcpy(all + E, jj, sizeof(int64_t) * E);
  qsort(all, 2 * (size_t)E, sizeof(int64_t), orc_cmp_i64);
  int nu = 0;
  for (int k = 0; k < 2 * E; k++)
    if (k == 0 || all[k] != all[k - 1]) uniq[nu++] = all[k];
  free(all);
  for (int e = 0; e < E; e++) {
    int lo = 0, hi = nu; /* searchsorted left */
    while (lo < hi) { int mid = (lo + hi) / 2; if (uniq[mid] < ii[e]) lo = mid + 1; else hi = mid; }
    ii_edge[e] = lo;
    lo = 0; hi = nu;
    while (lo < hi) { int mid = (lo + hi) / 2; if (uniq[mid] < jj[e]) lo = mid + 1; else hi = mid; }
    jj_edge[e] = lo;
  }
  return nu;
}

/* ------------------------------------------------------------------------------------------------
 * SparseBlock (gn_kernels.cu:57-159): assemble the 7(N-1) system in fp64 from fp32 blocks, drop rows/cols of
 * the fixed pose (index -1, :86,107), LLT, dx = -solve (:1209), zeros on factorisation failure (:147-150).
 * Eigen's SimplicialLLT<SparseMatrix<double>> is replaced by a dense fp64 Cholesky: same factorisation up
 * to the (AMD) elimination order, i.e. equal up to fp64 rounding.
 * ---------------------------------------------------------------------------------------------- */
ORC_API int orc_gn_solve(const float* Hs, const float* gs, const int64_t* ii_opt, const int64_t* jj_opt, int E,
                         int n_free, float* dx /*n_free,7*/, double* H_out /*optional dim*dim*/, double* b_out) {
  const int dim = 7 * n_free;
  double* A = (double*)calloc((size_t)dim * dim + 1, sizeof(double));
  double* b = (double*)calloc((size_t)dim + 1, sizeof(double));
  for (int blk = 0; blk < 4; blk++) {
    for (int e = 0; e < E; e++) {
      const int64_t i = (blk < 2) ? ii_opt[e] : jj_opt[e];
      const int64_t j = (blk % 2 == 0) ? ii_opt[e] : jj_opt[e];
      if (i >= 0 && j >= 0) {
        const float* M = Hs + ((int64_t)blk * E + e) * 49;
        for (int k = 0; k < 7; k++)
          for (int l = 0; l < 7; l++) A[(size_t)(7 * i + k) * dim + (7 * j + l)] += (double)M[k * 7 + l];
      }
    }
  }
  for (int blk = 0; blk < 2; blk++) {
    for (int e = 0; e < E; e++) {
      const int64_t i = blk == 0 ? ii_opt[e] : jj_opt[e];
      if (i >= 0)
        for (int k = 0; k < 7; k++) b[7 * i + k] += (double)gs[((int64_t)blk * E + e) * 7 + k];
    }
  }
  if (H_out) memcpy(H_out, A, sizeof(double) * (size_t)dim * dim);
  if (b_out) memcpy(b_out, b, sizeof(double) * (size_t)dim);
  int ok = 1;
  for (int c = 0; c < dim && ok; c++) { /* in-place lower Cholesky, column by column */
    double d = A[(size_t)c * dim + c];
    for (int k = 0; k < c; k++) d -= A[(size_t)c * dim + k] * A[(size_t)c * dim + k];
    if (!(d > 0.0)) { ok = 0; break; }
    d = sqrt(d);
    A[(size_t)c * dim + c] = d;
    for (int r = c + 1; r < dim; r++) {
      double s = A[(size_t)r * dim + c];
      for (int k = 0; k < c; k++) s -= A[(size_t)r * dim + k] * A[(size_t)c * dim + k];
      A[(size_t)r * dim + c] = s / d;
    }
  }
  if (ok) {
    for (int r = 0; r < dim; r++) { /* L y = b */
      double s = b[r];
      for (int k = 0; k < r; k++) s -= A[(size_t)r * dim + k] * b[k];
      b[r] = s / A[(size_t)r * dim + r];
    }
    for (int r = dim - 1; r >= 0; r--) { /* L^T x = y */
      double s = b[r];
      for (int k = r + 1; k < dim; k++) s -= A[(size_t)k * dim + r] * b[k];
      b[r] = s / A[(size_t)r * dim + r];
    }
    for (int r = 0; r < dim; r++) dx[r] = -(float)b[r];
  } else {
    for (int r = 0; r < dim; r++) dx[r] = 0.0f;
  }
  free(A);
  free(b);
  return ok ? 0 : 1;
}

/* ------------------------------------------------------------------------------------------------
 * Host GN loops gauss_newton_{points,rays,calib}_cuda (gn_kernels.cu:725-811, :1140-1228, :1546-1637).
 * Twc (num_kf,8) is updated in place; dx (num_kf-1,7) is the last update; returns iterations executed.
 * ---------------------------------------------------------------------------------------------- */
ORC_API int orc_gauss_newton(int kind, const float* params, int robust, int accum, float* Twc, const float* Xs,
                             const float* Cs, const int64_t* ii, const int64_t* jj, const int64_t* idx,
                             const uint8_t* valid_match, const float* Q, int num_kf, int npts, int E, float C_thresh,
                             float Q_thresh, int max_iter, float delta_thresh, float* dx /*num_kf-1,7*/) {
  const int num_fix = 1; /* :741,1157,1566 */
  int64_t* uniq = (int64_t*)malloc(sizeof(int64_t) * 2 * (size_t)(E + 1));
  int64_t* ie = (int64_t*)malloc(sizeof(int64_t) * (size_t)(E + 1));
  int64_t* je = (int64_t*)malloc(sizeof(int64_t) * (size_t)(E + 1));
  int64_t* io = (int64_t*)malloc(sizeof(int64_t) * (size_t)(E + 1));
  int64_t* jo = (int64_t*)malloc(sizeof(int64_t) * (size_t)(E + 1));
  orc_index_prep(ii, jj, E, uniq, ie, je);
  for (int e = 0; e < E; e++) { io[e] = ie[e] - num_fix; jo[e] = je[e] - num_fix; }
  float* Hs = (float*)calloc((size_t)4 * E * 49 + 1, sizeof(float));
  float* gs = (float*)calloc((size_t)2 * E * 7 + 1, sizeof(float));
  int it = 0;
  for (; it < max_iter;) {
    orc_gn_edge_blocks(kind, params, robust, accum, Twc, Xs, Cs, ie, je, idx, valid_match, Q, num_kf, npts, E,
                       C_thresh, Q_thresh, Hs, gs);
    orc_gn_solve(Hs, gs, io, jo, E, num_kf - num_fix, dx, NULL, NULL);
    orc_pose_retr(Twc, dx, num_kf, num_fix);
    it++;
    double nrm = 0.0; /* at::linalg_norm on a float tensor: float result */
    for (int k = 0; k < 7 * (num_kf - num_fix); k++) nrm += (double)dx[k] * (double)dx[k];
    if ((float)sqrt(nrm) < delta_thresh) break;
  }
  free(uniq); free(ie); free(je); free(io); free(jo); free(Hs); free(gs);
  return it;
}

/* explicit thread count for the timed CPU baseline (a launcher may have exported OMP_NUM_THREADS=1) */
ORC_API void orc_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

ORC_API int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
