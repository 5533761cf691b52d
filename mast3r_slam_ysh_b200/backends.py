"""Host-side mirror of the reference's `mast3r_slam_backends` pybind module (gn.cpp:116-123).

Same five functions, same positional signatures, tensor layouts, in-place update of `Twc` and list return values, so
`mast3r_slam/matching.py`, `global_opt.py` and `tracker.py` of the reference can call it unchanged
(`import mast3r_slam_backends` resolves to the shim at the repo root).  All compute happens in
libm3r_backends_b200.so through the C ABI of include/m3r_backends.h, on torch's CURRENT stream; torch is used only
for device memory and stream handles.  No CPU fallback exists: CPU tensors raise.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import GnParams

__all__ = ["gauss_newton_points", "gauss_newton_rays", "gauss_newton_calib", "iter_proj", "refine_matches",
           "gn_edge_blocks", "pose_retr", "launch_count"]


def _check(t, name, dtype=None, ndim=None):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if not t.is_contiguous():  # CHECK_CONTIGUOUS, gn.h:5
        raise RuntimeError(f"{name} must be contiguous")
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (the B200 backend has no CPU path)")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"{name}: expected dtype {dtype}, got {t.dtype}")
    if ndim is not None and t.dim() != ndim:
        raise RuntimeError(f"{name}: expected {ndim} dimensions, got {t.dim()}")


def _same_device(ref, *ts):
    for t in ts:
        if t.device != ref.device:
            raise RuntimeError("all tensors must be on the same CUDA device")


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)


def launch_count():
    return int(_lib.load().m3r_launch_count())


# ------------------------------------------------------------------------------------------------------ matching
def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh):
    """reference: iter_proj (gn.cpp:84-99; matching_kernels.cu:279-316). Returns [p_new (b,n,2) f32, converged (b,n) bool]."""
    _check(rays_img_with_grad, "rays_img_with_grad", torch.float32, 4)
    _check(pts_3d_norm, "pts_3d_norm", torch.float32, 3)
    _check(p_init, "p_init", torch.float32, 3)
    _same_device(p_init, rays_img_with_grad, pts_3d_norm)
    b, h, w, c = rays_img_with_grad.shape
    if c != 9:
        raise RuntimeError("rays_img_with_grad must have 9 channels (ray, d/du, d/dv)")
    n = p_init.shape[1]
    if p_init.shape[0] != b or p_init.shape[2] != 2 or tuple(pts_3d_norm.shape) != (b, n, 3):
        raise RuntimeError("iter_proj: inconsistent shapes")
    p_new = torch.zeros((b, n, 2), dtype=torch.float32, device=p_init.device)
    converged = torch.zeros((b, n), dtype=torch.bool, device=p_init.device)
    with torch.cuda.device(p_init.device):
        rc = _lib.load().m3r_iter_proj(_ptr(rays_img_with_grad), _ptr(pts_3d_norm), _ptr(p_init), _ptr(p_new),
                                       _ptr(converged), b, h, w, n, int(max_iter), float(lambda_init),
                                       float(cost_thresh), _stream(p_init.device))
    _lib.check(rc, "iter_proj")
    return [p_new, converged]


def refine_matches(D11, D21, p1, radius, dilation_max):
    """reference: refine_matches (gn.cpp:101-114; matching_kernels.cu:84-116). Returns [p1_new (b,n,2) int64]."""
    _check(D11, "D11", None, 4)
    _check(D21, "D21", None, 3)
    _check(p1, "p1", torch.int64, 3)
    _same_device(p1, D11, D21)
    if D11.dtype != D21.dtype or D11.dtype not in (torch.float16, torch.float32):
        raise RuntimeError("refine_matches: D11/D21 must both be float16 (or float32)")
    b, h, w, fdim = D11.shape
    n = p1.shape[1]
    if p1.shape[0] != b or p1.shape[2] != 2 or tuple(D21.shape) != (b, n, fdim):
        raise RuntimeError("refine_matches: inconsistent shapes")
    p1_new = torch.zeros((b, n, 2), dtype=torch.int64, device=p1.device)
    lib = _lib.load()
    fn = lib.m3r_refine_matches_f16 if D11.dtype == torch.float16 else lib.m3r_refine_matches_f32
    with torch.cuda.device(p1.device):
        rc = fn(_ptr(D11), _ptr(D21), _ptr(p1), _ptr(p1_new), b, h, w, n, fdim, int(radius), int(dilation_max),
                _stream(p1.device))
    _lib.check(rc, "refine_matches")
    return [p1_new]


# ------------------------------------------------------------------------------------------------------ Gauss-Newton
def _gn_check(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q):
    _check(Twc, "Twc", torch.float32, 2)
    _check(Xs, "Xs", torch.float32, 3)
    _check(Cs, "Cs", torch.float32)
    _check(ii, "ii", torch.int64, 1)
    _check(jj, "jj", torch.int64, 1)
    _check(idx_ii2jj, "idx_ii2jj", torch.int64, 2)
    _check(valid_match, "valid_match", torch.bool)
    _check(Q, "Q", torch.float32)
    _same_device(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q)
    num_kf, num_pts = Xs.shape[0], Xs.shape[1]
    E = ii.shape[0]
    if Twc.shape[0] != num_kf or Twc.shape[1] != 8 or Xs.shape[2] != 3:
        raise RuntimeError("gauss_newton: Twc must be (num_kf,8) and Xs (num_kf,N,3)")
    if Cs.numel() != num_kf * num_pts or jj.shape[0] != E or tuple(idx_ii2jj.shape) != (E, num_pts):
        raise RuntimeError("gauss_newton: inconsistent shapes")
    if valid_match.numel() != E * num_pts or Q.numel() != E * num_pts:
        raise RuntimeError("gauss_newton: valid_match/Q must have num_edges*num_pts elements")
    return num_kf, num_pts, E


def _gn_params(kind, sigma_a, sigma_b, C_thresh, Q_thresh, max_iter, delta_thresh, robust, K=None, height=0, width=0,
               pixel_border=0, z_eps=0.0):
    rb = {"huber": _lib.M3R_ROBUST_HUBER, "tukey": _lib.M3R_ROBUST_TUKEY}[robust]
    return GnParams(kind, rb, float(sigma_a), float(sigma_b), float(C_thresh), float(Q_thresh),
                    K.data_ptr() if K is not None else None, int(height), int(width), int(pixel_border), float(z_eps),
                    int(max_iter), float(delta_thresh))


def _gn_run(params, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, return_info=False):
    num_kf, num_pts, E = _gn_check(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q)
    dev = Twc.device
    lib = _lib.load()
    dx = torch.zeros((max(num_kf - 1, 0), 7), dtype=torch.float32, device=dev)
    info = torch.zeros(4, dtype=torch.int32, device=dev)
    wsb = lib.m3r_gn_workspace_bytes(num_kf, num_pts, E)
    ws = torch.empty(wsb + 256, dtype=torch.uint8, device=dev)
    off = (-ws.data_ptr()) % 256
    with torch.cuda.device(dev):
        rc = lib.m3r_gauss_newton(C.byref(params), _ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(ii), _ptr(jj), _ptr(idx_ii2jj),
                                  _ptr(valid_match), _ptr(Q), num_kf, num_pts, E, _ptr(dx), _ptr(info),
                                  C.c_void_p(ws.data_ptr() + off), wsb, _stream(dev))
    _lib.check(rc, "gauss_newton")
    ws.record_stream(torch.cuda.current_stream(dev))
    return ([dx], info) if return_info else [dx]


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_point, C_thresh, Q_thresh, max_iter,
                        delta_thresh, *, robust="huber", return_info=False):
    """reference: gauss_newton_points (gn.cpp:3-26; gn_kernels.cu:725-811). Updates Twc in place; returns [dx]."""
    p = _gn_params(_lib.M3R_GN_POINTS, sigma_point, 1.0, C_thresh, Q_thresh, max_iter, delta_thresh, robust)
    return _gn_run(p, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, return_info)


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray, sigma_dist, C_thresh, Q_thresh,
                      max_iter, delta_thresh, *, robust="huber", return_info=False):
    """reference: gauss_newton_rays (gn.cpp:28-52; gn_kernels.cu:1140-1228). Updates Twc in place; returns [dx]."""
    p = _gn_params(_lib.M3R_GN_RAYS, sigma_ray, sigma_dist, C_thresh, Q_thresh, max_iter, delta_thresh, robust)
    return _gn_run(p, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, return_info)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, height, width, pixel_border, z_eps,
                       sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter, delta_thresh, *, robust="huber",
                       return_info=False):
    """reference: gauss_newton_calib (gn.cpp:54-82; gn_kernels.cu:1546-1637). Updates Twc in place; returns [dx]."""
    _check(K, "K", torch.float32, 2)
    if tuple(K.shape) != (3, 3):
        raise RuntimeError("K must be (3,3)")
    p = _gn_params(_lib.M3R_GN_CALIB, sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter, delta_thresh, robust, K,
                   height, width, pixel_border, z_eps)
    return _gn_run(p, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, return_info)


# ------------------------------------------------------------------------------------------------------ diagnostics
def gn_edge_blocks(kind, Twc, Xs, Cs, ii_edge, jj_edge, idx_ii2jj, valid_match, Q, C_thresh, Q_thresh, robust="huber",
                   **kw):
    """One evaluation of the per-edge kernel in the reference's layout: Hs (4,E,7,7), gs (2,E,7)
    (ray_align / calib_proj / point_align epilogue, gn_kernels.cu:1097-1137). ii_edge/jj_edge index Twc rows."""
    num_kf, num_pts, E = _gn_check(Twc, Xs, Cs, ii_edge, jj_edge, idx_ii2jj, valid_match, Q)
    if kind == "points":
        p = _gn_params(_lib.M3R_GN_POINTS, kw["sigma_point"], 1.0, C_thresh, Q_thresh, 1, 0.0, robust)
    elif kind == "rays":
        p = _gn_params(_lib.M3R_GN_RAYS, kw["sigma_ray"], kw["sigma_dist"], C_thresh, Q_thresh, 1, 0.0, robust)
    else:
        p = _gn_params(_lib.M3R_GN_CALIB, kw["sigma_pixel"], kw["sigma_depth"], C_thresh, Q_thresh, 1, 0.0, robust,
                       kw["K"], kw["height"], kw["width"], kw["pixel_border"], kw["z_eps"])
    dev = Twc.device
    lib = _lib.load()
    Hs = torch.zeros((4, E, 7, 7), dtype=torch.float32, device=dev)
    gs = torch.zeros((2, E, 7), dtype=torch.float32, device=dev)
    wsb = lib.m3r_gn_workspace_bytes(num_kf, num_pts, E)
    ws = torch.empty(wsb + 256, dtype=torch.uint8, device=dev)
    off = (-ws.data_ptr()) % 256
    with torch.cuda.device(dev):
        rc = lib.m3r_gn_edge_blocks(C.byref(p), _ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(ii_edge), _ptr(jj_edge),
                                    _ptr(idx_ii2jj), _ptr(valid_match), _ptr(Q), num_kf, num_pts, E, _ptr(Hs),
                                    _ptr(gs), C.c_void_p(ws.data_ptr() + off), wsb, _stream(dev))
    _lib.check(rc, "gn_edge_blocks")
    ws.record_stream(torch.cuda.current_stream(dev))
    return Hs, gs


def pose_retr(poses, dx, num_fix=1):
    """pose_retr_kernel (gn_kernels.cu:415-453): in place on poses (num_poses,8)."""
    _check(poses, "poses", torch.float32, 2)
    _check(dx, "dx", torch.float32, 2)
    with torch.cuda.device(poses.device):
        rc = _lib.load().m3r_pose_retr(_ptr(poses), _ptr(dx), poses.shape[0], int(num_fix), _stream(poses.device))
    _lib.check(rc, "pose_retr")
    return poses
