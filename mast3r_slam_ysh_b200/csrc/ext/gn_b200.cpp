// Compiled drop-in for the reference's pybind module (backend/src/gn.cpp:116-123, declarations backend/include/gn.h:11-117):
// a thin torch C++ extension that keeps the five function names, positional signatures, tensor layouts, the in-place
// update of Twc and the std::vector<torch::Tensor> return values, and forwards to the C ABI of include/m3r_backends.h
// (libm3r_backends_b200.so) on torch's current CUDA stream.  Built in-tree by __graft_entry__.build() (g++, no nvcc: the
// kernels live in the C-ABI library) as mast3r_slam_ysh_b200/mast3r_slam_backends_b200*.so; the repo-root shim
// mast3r_slam_backends.py re-exports it.  No CPU path: CPU tensors fail the CUDA check below.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include <vector>

#include "m3r_backends.h"

namespace {

// gn.h:5 -- CHECK_CONTIGUOUS; the device check is the one the reference defines but never applies (gn.h:6)
#define B200_CHECK_INPUT(x)                                                    \
  TORCH_CHECK((x).is_contiguous(), #x " must be contiguous");                  \
  TORCH_CHECK((x).is_cuda(), #x " must be a CUDA tensor (the B200 backend has no CPU path)")

void check_rc(int rc, const char* what) {
  TORCH_CHECK(rc == 0, what, " failed (code ", rc, "): ", m3r_last_error_string());
}

void* stream_of(const torch::Tensor& t) { return (void*)at::cuda::getCurrentCUDAStream(t.get_device()).stream(); }

struct GnCall {
  m3r_gn_params p{};
  torch::Tensor Twc, Xs, Cs, ii, jj, idx, valid, Q;
};

std::vector<torch::Tensor> run_gn(GnCall& c, bool want_info) {
  B200_CHECK_INPUT(c.Twc); B200_CHECK_INPUT(c.Xs); B200_CHECK_INPUT(c.Cs); B200_CHECK_INPUT(c.ii); B200_CHECK_INPUT(c.jj);
  B200_CHECK_INPUT(c.idx); B200_CHECK_INPUT(c.valid); B200_CHECK_INPUT(c.Q);
  TORCH_CHECK(c.Twc.scalar_type() == torch::kFloat32 && c.Xs.scalar_type() == torch::kFloat32 &&
                  c.Cs.scalar_type() == torch::kFloat32 && c.Q.scalar_type() == torch::kFloat32,
              "gauss_newton: Twc/Xs/Cs/Q must be float32");
  TORCH_CHECK(c.ii.scalar_type() == torch::kInt64 && c.jj.scalar_type() == torch::kInt64 &&
                  c.idx.scalar_type() == torch::kInt64, "gauss_newton: ii/jj/idx_ii2jj must be int64");
  TORCH_CHECK(c.valid.scalar_type() == torch::kBool, "gauss_newton: valid_match must be bool");
  TORCH_CHECK(c.Twc.dim() == 2 && c.Twc.size(1) == 8 && c.Xs.dim() == 3 && c.Xs.size(2) == 3 &&
                  c.Xs.size(0) == c.Twc.size(0), "gauss_newton: Twc must be (num_kf,8) and Xs (num_kf,N,3)");
  const int num_kf = (int)c.Xs.size(0), num_pts = (int)c.Xs.size(1), num_edges = (int)c.ii.size(0);
  TORCH_CHECK(c.Cs.numel() == (int64_t)num_kf * num_pts && c.jj.size(0) == num_edges &&
                  c.idx.numel() == (int64_t)num_edges * num_pts && c.valid.numel() == c.idx.numel() &&
                  c.Q.numel() == c.idx.numel(), "gauss_newton: inconsistent shapes");
  const c10::cuda::CUDAGuard guard(c.Twc.device());
  auto dx = torch::zeros({num_kf > 0 ? num_kf - 1 : 0, 7}, c.Twc.options());
  auto info = torch::zeros({4}, c.Twc.options().dtype(torch::kInt32));
  const size_t wsb = m3r_gn_workspace_bytes(num_kf, num_pts, num_edges);
  auto ws = torch::empty({(int64_t)wsb + 256}, c.Twc.options().dtype(torch::kUInt8));
  char* wsp = (char*)ws.data_ptr();
  wsp += (256 - (uintptr_t)wsp % 256) % 256;
  auto ptr = [](const torch::Tensor& t) -> void* { return t.numel() > 0 ? t.data_ptr() : nullptr; };
  check_rc(m3r_gauss_newton(&c.p, (float*)ptr(c.Twc), (const float*)ptr(c.Xs), (const float*)ptr(c.Cs),
                            (const int64_t*)ptr(c.ii), (const int64_t*)ptr(c.jj), (const int64_t*)ptr(c.idx),
                            (const uint8_t*)ptr(c.valid), (const float*)ptr(c.Q), num_kf, num_pts, num_edges,
                            (float*)ptr(dx), (int32_t*)info.data_ptr(), wsp, wsb, stream_of(c.Twc)),
           "gauss_newton");
  // the workspace is only referenced by enqueued work: keep the allocator from handing it out before that has run
  c10::cuda::CUDACachingAllocator::recordStream(ws.storage().data_ptr(), at::cuda::getCurrentCUDAStream(c.Twc.get_device()));
  if (want_info) return {dx, info};
  return {dx};
}

}  // namespace

// gn.h:22-31
std::vector<torch::Tensor> gauss_newton_points(torch::Tensor Twc, torch::Tensor Xs, torch::Tensor Cs, torch::Tensor ii,
                                               torch::Tensor jj, torch::Tensor idx_ii2jj, torch::Tensor valid_match,
                                               torch::Tensor Q, const float sigma_point, const float C_thresh,
                                               const float Q_thresh, const int max_iter, const float delta_thresh) {
  GnCall c{{}, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q};
  c.p.kind = M3R_GN_POINTS; c.p.robust = M3R_ROBUST_HUBER; c.p.sigma_a = sigma_point; c.p.sigma_b = 1.0f;
  c.p.C_thresh = C_thresh; c.p.Q_thresh = Q_thresh; c.p.max_iter = max_iter; c.p.delta_thresh = delta_thresh;
  return run_gn(c, false);
}

// gn.h:45-55
std::vector<torch::Tensor> gauss_newton_rays(torch::Tensor Twc, torch::Tensor Xs, torch::Tensor Cs, torch::Tensor ii,
                                             torch::Tensor jj, torch::Tensor idx_ii2jj, torch::Tensor valid_match,
                                             torch::Tensor Q, const float sigma_ray, const float sigma_dist,
                                             const float C_thresh, const float Q_thresh, const int max_iter,
                                             const float delta_thresh) {
  GnCall c{{}, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q};
  c.p.kind = M3R_GN_RAYS; c.p.robust = M3R_ROBUST_HUBER; c.p.sigma_a = sigma_ray; c.p.sigma_b = sigma_dist;
  c.p.C_thresh = C_thresh; c.p.Q_thresh = Q_thresh; c.p.max_iter = max_iter; c.p.delta_thresh = delta_thresh;
  return run_gn(c, false);
}

// gn.h:74-87
std::vector<torch::Tensor> gauss_newton_calib(torch::Tensor Twc, torch::Tensor Xs, torch::Tensor Cs, torch::Tensor K,
                                              torch::Tensor ii, torch::Tensor jj, torch::Tensor idx_ii2jj,
                                              torch::Tensor valid_match, torch::Tensor Q, const int height,
                                              const int width, const int pixel_border, const float z_eps,
                                              const float sigma_pixel, const float sigma_depth, const float C_thresh,
                                              const float Q_thresh, const int max_iter, const float delta_thresh) {
  B200_CHECK_INPUT(K);
  TORCH_CHECK(K.scalar_type() == torch::kFloat32 && K.dim() == 2 && K.size(0) == 3 && K.size(1) == 3, "K must be (3,3) float32");
  GnCall c{{}, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q};
  c.p.kind = M3R_GN_CALIB; c.p.robust = M3R_ROBUST_HUBER; c.p.sigma_a = sigma_pixel; c.p.sigma_b = sigma_depth;
  c.p.C_thresh = C_thresh; c.p.Q_thresh = Q_thresh; c.p.K = K.data_ptr<float>(); c.p.height = height; c.p.width = width;
  c.p.pixel_border = pixel_border; c.p.z_eps = z_eps; c.p.max_iter = max_iter; c.p.delta_thresh = delta_thresh;
  return run_gn(c, false);
}

// gn.h:89-95
std::vector<torch::Tensor> iter_proj(torch::Tensor rays_img_with_grad, torch::Tensor pts_3d_norm, torch::Tensor p_init,
                                     const int max_iter, const float lambda_init, const float cost_thresh) {
  B200_CHECK_INPUT(rays_img_with_grad); B200_CHECK_INPUT(pts_3d_norm); B200_CHECK_INPUT(p_init);
  TORCH_CHECK(rays_img_with_grad.scalar_type() == torch::kFloat32 && pts_3d_norm.scalar_type() == torch::kFloat32 &&
                  p_init.scalar_type() == torch::kFloat32, "iter_proj: float32 tensors expected");
  TORCH_CHECK(rays_img_with_grad.dim() == 4 && rays_img_with_grad.size(3) == 9, "rays_img_with_grad must be (b,h,w,9)");
  TORCH_CHECK(p_init.dim() == 3 && p_init.size(2) == 2 && pts_3d_norm.dim() == 3 && pts_3d_norm.size(2) == 3 &&
                  p_init.size(0) == rays_img_with_grad.size(0) && pts_3d_norm.size(0) == p_init.size(0) &&
                  pts_3d_norm.size(1) == p_init.size(1), "iter_proj: inconsistent shapes");
  const int b = (int)p_init.size(0), n = (int)p_init.size(1), h = (int)rays_img_with_grad.size(1),
            w = (int)rays_img_with_grad.size(2);
  const c10::cuda::CUDAGuard guard(p_init.device());
  auto p_new = torch::zeros({b, n, 2}, p_init.options());
  auto conv = torch::zeros({b, n}, p_init.options().dtype(torch::kBool));
  check_rc(m3r_iter_proj(rays_img_with_grad.data_ptr<float>(), pts_3d_norm.data_ptr<float>(), p_init.data_ptr<float>(),
                         p_new.data_ptr<float>(), (uint8_t*)conv.data_ptr<bool>(), b, h, w, n, max_iter, lambda_init,
                         cost_thresh, stream_of(p_init)), "iter_proj");
  return {p_new, conv};
}

// gn.h:112-117
std::vector<torch::Tensor> refine_matches(torch::Tensor D11, torch::Tensor D21, torch::Tensor p1, const int radius,
                                          const int dilation_max) {
  B200_CHECK_INPUT(D11); B200_CHECK_INPUT(D21); B200_CHECK_INPUT(p1);
  TORCH_CHECK(p1.scalar_type() == torch::kInt64, "p1 must be int64");
  TORCH_CHECK(D11.scalar_type() == D21.scalar_type() &&
                  (D11.scalar_type() == torch::kFloat16 || D11.scalar_type() == torch::kFloat32),
              "refine_matches: D11/D21 must both be float16 (or float32)");
  TORCH_CHECK(D11.dim() == 4 && D21.dim() == 3 && p1.dim() == 3 && p1.size(2) == 2 && p1.size(0) == D11.size(0) &&
                  D21.size(0) == D11.size(0) && D21.size(1) == p1.size(1) && D21.size(2) == D11.size(3),
              "refine_matches: inconsistent shapes");
  const int b = (int)D11.size(0), h = (int)D11.size(1), w = (int)D11.size(2), fdim = (int)D11.size(3), n = (int)p1.size(1);
  const c10::cuda::CUDAGuard guard(p1.device());
  auto p1_new = torch::zeros({b, n, 2}, p1.options());
  if (D11.scalar_type() == torch::kFloat16)
    check_rc(m3r_refine_matches_f16((const uint16_t*)D11.data_ptr(), (const uint16_t*)D21.data_ptr(),
                                    p1.data_ptr<int64_t>(), p1_new.data_ptr<int64_t>(), b, h, w, n, fdim, radius,
                                    dilation_max, stream_of(p1)), "refine_matches");
  else
    check_rc(m3r_refine_matches_f32(D11.data_ptr<float>(), D21.data_ptr<float>(), p1.data_ptr<int64_t>(),
                                    p1_new.data_ptr<int64_t>(), b, h, w, n, fdim, radius, dilation_max, stream_of(p1)),
             "refine_matches");
  return {p1_new};
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {  // the function table of gn.cpp:116-123
  m.def("gauss_newton_points", &gauss_newton_points, "point alignment");
  m.def("gauss_newton_rays", &gauss_newton_rays, "ray alignment");
  m.def("gauss_newton_calib", &gauss_newton_calib, "calibrated reprojection");
  m.def("iter_proj", &iter_proj, "iterative projective matching");
  m.def("refine_matches", &refine_matches, "descriptor-window match refinement");
  m.def("launch_count", []() { return (uint64_t)m3r_launch_count(); }, "kernels launched by the C-ABI library so far");
}
