// Sim(3) retraction shared by pose_retr_kernel and the fused solver kernel.
#pragma once
#include "gn_internal.cuh"

namespace m3r {

// ============================================================================================== Sim(3) retraction
// expSO3 / expSim3 / retrSim3 (gn_kernels.cu:299-413): T <- Exp(xi) o T with xi = (tau, phi, sigma); the
// quaternion is not renormalised.
__device__ __forceinline__ void quat_comp_d(const float* qi, const float* qj, float* out) {
  out[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  out[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  out[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  out[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}

__device__ __forceinline__ void act_so3_d(const float* q, const float* X, float* Y) {
  const float uv0 = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  const float uv1 = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  const float uv2 = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  const float y0 = X[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
  const float y1 = X[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
  const float y2 = X[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
  Y[0] = y0; Y[1] = y1; Y[2] = y2;
}

__device__ __forceinline__ void retr_sim3(const float* xi, float* pose) {
  const float kEps = 1e-6f;
  float tau[3] = {xi[0], xi[1], xi[2]};
  const float phi[3] = {xi[3], xi[4], xi[5]};
  const float sigma = xi[6];
  const float scale = expf(sigma);
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float theta = sqrtf(theta_sq);
  float dq[4];
  {
    float imag, real;
    if (theta_sq < kEps) {
      const float theta_p4 = theta_sq * theta_sq;
      imag = 0.5f - (1.0f / 48.0f) * theta_sq + (1.0f / 3840.0f) * theta_p4;
      real = 1.0f - (1.0f / 8.0f) * theta_sq + (1.0f / 384.0f) * theta_p4;
    } else {
      imag = sinf(0.5f * theta) / theta;
      real = cosf(0.5f * theta);
    }
    dq[0] = imag * phi[0]; dq[1] = imag * phi[1]; dq[2] = imag * phi[2]; dq[3] = real;
  }
  float A, B, C;
  if (fabsf(sigma) < kEps) {
    C = 1.0f;
    if (fabsf(theta) < kEps) {
      A = 0.5f;
      B = 1.0f / 6.0f;
    } else {
      A = (1.0f - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - 1.0f) / sigma;
    if (fabsf(theta) < kEps) {
      const float sigma_sq = sigma * sigma;
      A = ((sigma - 1.0f) * scale + 1.0f) / sigma_sq;
      B = (scale * 0.5f * sigma_sq + scale - 1.0f - sigma * scale) / (sigma_sq * sigma);
    } else {
      const float a = scale * sinf(theta);
      const float b = scale * cosf(theta);
      const float c = theta_sq + sigma * sigma;
      A = (a * sigma + (1.0f - b) * theta) / (theta * c);
      B = (C - ((b - 1.0f) * sigma + a * theta) / c) / theta_sq;
    }
  }
  float dt[3] = {C * tau[0], C * tau[1], C * tau[2]};
  float cr[3] = {phi[1] * tau[2] - phi[2] * tau[1], phi[2] * tau[0] - phi[0] * tau[2], phi[0] * tau[1] - phi[1] * tau[0]};
  dt[0] += A * cr[0]; dt[1] += A * cr[1]; dt[2] += A * cr[2];
  float cr2[3] = {phi[1] * cr[2] - phi[2] * cr[1], phi[2] * cr[0] - phi[0] * cr[2], phi[0] * cr[1] - phi[1] * cr[0]};
  dt[0] += B * cr2[0]; dt[1] += B * cr2[1]; dt[2] += B * cr2[2];

  const float t[3] = {pose[0], pose[1], pose[2]};
  const float q[4] = {pose[3], pose[4], pose[5], pose[6]};
  const float s = pose[7];
  float q1[4], t1[3];
  quat_comp_d(dq, q, q1);
  act_so3_d(dq, t, t1);
  pose[0] = t1[0] * scale + dt[0];
  pose[1] = t1[1] * scale + dt[1];
  pose[2] = t1[2] * scale + dt[2];
  pose[3] = q1[0]; pose[4] = q1[1]; pose[5] = q1[2]; pose[6] = q1[3];
  pose[7] = scale * s;
}

}  // namespace m3r
