// Small element-wise geometry passes of the graph solve's callers.
//   m3r_constrain_points_to_ray : geometry.constrain_points_to_ray (mast3r_slam/geometry.py:37-42) = backproject
//   (:111-120) of the pixel grid (get_pixel_coords, :123-129) with the point's own depth: X = z * [(u-cx)/fx, (v-cy)/fy, 1].
//   Called per solve on all keyframe pointmaps (global_opt.py:172) and per tracked frame (tracker.py:141-142).
// The float sequence is the reference's (subtract, IEEE divide, multiply; no contraction possible), so the result is
// bit-identical to the torch ops.  HBM-bound: 12 B read + 12 B written per point, in place if out == Xs.
//   m3r_update_pointmap : Frame.update_pointmap (mast3r_slam/frame.py:41-105) for the running filters
//   "weighted_pointmap" (config default) and "indep_conf", in place, optionally fused with the Sim(3) action the tracker
//   applies first (Xkk = T_CkCf.act(Xkf), tracker.py:96-97): 3-4 eager launches and two temporaries -> one pass.
#include "gn_sim3.cuh"

namespace m3r {

// mode 0: X = (C X + Cn Xn) / (C + Cn), C += Cn (frame.py:73-76); mode 1: where Cn > C take (Xn, Cn) (frame.py:68-72).
// The float sequence is the reference's (separate multiplies, add, IEEE divide), so without a pose the result is
// bit-identical to the torch ops.
__global__ void __launch_bounds__(256) update_pointmap_kernel(float* __restrict__ X, float* __restrict__ Cc,
                                                              const float* __restrict__ Xn, const float* __restrict__ Cn,
                                                              const float* __restrict__ T, long long n, int mode) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float p[3] = {Xn[3 * i], Xn[3 * i + 1], Xn[3 * i + 2]};
  if (T != nullptr) {  // lietorch Sim3.act: s * (q * p) + t
    const float t[3] = {__ldg(T), __ldg(T + 1), __ldg(T + 2)};
    const float q[4] = {__ldg(T + 3), __ldg(T + 4), __ldg(T + 5), __ldg(T + 6)};
    const float sc = __ldg(T + 7);
    float r[3];
    act_so3_d(q, p, r);
    p[0] = sc * r[0] + t[0];
    p[1] = sc * r[1] + t[1];
    p[2] = sc * r[2] + t[2];
  }
  const float c = Cc[i], cn = Cn[i];
  if (mode == 0) {
    const float den = __fadd_rn(c, cn);
#pragma unroll
    for (int k = 0; k < 3; k++)
      X[3 * i + k] = __fdiv_rn(__fadd_rn(__fmul_rn(c, X[3 * i + k]), __fmul_rn(cn, p[k])), den);
    Cc[i] = den;
  } else if (mode == 2) {
    // weighted_spherical (frame.py:76-101): confidence-weighted mean of (r, phi, theta) = (|P|, atan2(y, x), acos(z / r)),
    // back to cartesian.  atan2f/acosf/sinf/cosf are the device libm versions: agreement with torch's CUDA ops to a few ulp
    // of the angles (tests/test_gpu_geometry.py), not bit for bit.
    const float den = __fadd_rn(c, cn);
    const float x0 = X[3 * i], y0 = X[3 * i + 1], z0 = X[3 * i + 2];
    const float r0 = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(x0, x0), __fmul_rn(y0, y0)), __fmul_rn(z0, z0)));
    const float r1 = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(p[0], p[0]), __fmul_rn(p[1], p[1])), __fmul_rn(p[2], p[2])));
    const float ph0 = atan2f(y0, x0), ph1 = atan2f(p[1], p[0]);
    const float th0 = acosf(__fdiv_rn(z0, r0)), th1 = acosf(__fdiv_rn(p[2], r1));
    const float r = __fdiv_rn(__fadd_rn(__fmul_rn(c, r0), __fmul_rn(cn, r1)), den);
    const float ph = __fdiv_rn(__fadd_rn(__fmul_rn(c, ph0), __fmul_rn(cn, ph1)), den);
    const float th = __fdiv_rn(__fadd_rn(__fmul_rn(c, th0), __fmul_rn(cn, th1)), den);
    const float st = sinf(th);
    X[3 * i] = __fmul_rn(__fmul_rn(r, st), cosf(ph));
    X[3 * i + 1] = __fmul_rn(__fmul_rn(r, st), sinf(ph));
    X[3 * i + 2] = __fmul_rn(r, cosf(th));
    Cc[i] = den;
  } else if (cn > c) {
    X[3 * i] = p[0];
    X[3 * i + 1] = p[1];
    X[3 * i + 2] = p[2];
    Cc[i] = cn;
  }
}

__global__ void __launch_bounds__(256) constrain_points_kernel(const float* __restrict__ Xs, const float* __restrict__ K,
                                                               float* __restrict__ out, long long total, int h, int w) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const float fx = __ldg(K + 0), cx = __ldg(K + 2), fy = __ldg(K + 4), cy = __ldg(K + 5);
  const int pix = (int)(i % ((long long)h * w));
  const int v = pix / w, u = pix - v * w;
  const float z = Xs[3 * i + 2];
  const float tx = __fdiv_rn(__fsub_rn((float)u, cx), fx);
  const float ty = __fdiv_rn(__fsub_rn((float)v, cy), fy);
  out[3 * i + 0] = __fmul_rn(z, tx);
  out[3 * i + 1] = __fmul_rn(z, ty);
  out[3 * i + 2] = z;  // z * 1.0
}

}  // namespace m3r

using namespace m3r;

extern "C" int m3r_constrain_points_to_ray(const float* Xs, const float* K, float* out, long long num_frames, int h,
                                           int w, void* stream) {
  M3R_REQUIRE(num_frames >= 0 && h >= 0 && w >= 0, "constrain_points_to_ray: negative size");
  const long long total = num_frames * h * w;
  if (total == 0) return M3R_OK;
  M3R_REQUIRE(Xs && K && out, "constrain_points_to_ray: null pointer");
  M3R_REQUIRE((long long)h * w < (1LL << 31), "constrain_points_to_ray: image too large");
  constrain_points_kernel<<<(unsigned)ceil_div_ll(total, 256), 256, 0, (cudaStream_t)stream>>>(Xs, K, out, total, h, w);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

extern "C" int m3r_update_pointmap(float* X_canon, float* C, const float* X_new, const float* C_new, const float* T,
                                   long long n, int mode, void* stream) {
  M3R_REQUIRE(n >= 0, "update_pointmap: negative size");
  M3R_REQUIRE(mode >= 0 && mode <= 2,
              "update_pointmap: mode must be 0 (weighted_pointmap), 1 (indep_conf) or 2 (weighted_spherical)");
  if (n == 0) return M3R_OK;
  M3R_REQUIRE(X_canon && C && X_new && C_new, "update_pointmap: null pointer");
  update_pointmap_kernel<<<(unsigned)ceil_div_ll(n, 256), 256, 0, (cudaStream_t)stream>>>(X_canon, C, X_new, C_new, T, n,
                                                                                         mode);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}
