// Small element-wise geometry passes of the graph solve's callers.
//   m3r_constrain_points_to_ray : geometry.constrain_points_to_ray (mast3r_slam/geometry.py:37-42) = backproject
//   (:111-120) of the pixel grid (get_pixel_coords, :123-129) with the point's own depth: X = z * [(u-cx)/fx, (v-cy)/fy, 1].
//   Called per solve on all keyframe pointmaps (global_opt.py:172) and per tracked frame (tracker.py:141-142).
// The float sequence is the reference's (subtract, IEEE divide, multiply; no contraction possible), so the result is
// bit-identical to the torch ops.  HBM-bound: 12 B read + 12 B written per point, in place if out == Xs.
#include "m3r_common.cuh"

namespace m3r {

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
