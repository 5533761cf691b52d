"""TEST INFRASTRUCTURE ONLY -- ctypes front end of the CPU oracle (oracle/m3r_oracle.c).

Mirrors the five `mast3r_slam_backends` entry points of the reference (gn.cpp:116-123) on numpy arrays so the
parity tests read like calls of the reference.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product package must never do so.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libm3r_oracle.so")
_lib = None

KIND_POINTS, KIND_RAYS, KIND_CALIB = 0, 1, 2
ROBUST_HUBER, ROBUST_TUKEY = 0, 1


def build(force=False):
    src = os.path.join(_HERE, "m3r_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.orc_gauss_newton.restype = C.c_int
        _lib.orc_index_prep.restype = C.c_int
        _lib.orc_gn_solve.restype = C.c_int
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    a = np.ascontiguousarray(a, dtype=dt)
    return a


def num_threads():
    return int(lib().orc_num_threads())


def set_num_threads(n):
    """OpenMP threads of the oracle for the timed CPU baseline (torchrun exports OMP_NUM_THREADS=1)."""
    lib().orc_set_num_threads(C.c_int(int(n)))
    return num_threads()


def f32_to_f16_bits(x):
    x = _c(x, np.float32)
    out = np.empty(x.shape, np.uint16)
    lib().orc_f32_to_f16(_p(x), _p(out), C.c_int64(x.size))
    return out


def f16_bits_to_f32(x):
    x = _c(x, np.uint16)
    out = np.empty(x.shape, np.float32)
    lib().orc_f16_to_f32(_p(x), _p(out), C.c_int64(x.size))
    return out


def robust_weights(r, robust=ROBUST_HUBER):
    """IRLS weight of the GN kernels: Huber k=1.345 (gn_kernels.cu:172-175) or Tukey t=4.6851
    (nonlinear_optimizer.py:36-42)."""
    r = _c(r, np.float32)
    out = np.empty(r.shape, np.float32)
    lib().orc_robust_weights(_p(r), C.c_int64(r.size), C.c_int(int(robust)), _p(out))
    return out


def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh):
    """reference: iter_proj (gn.cpp:84-99) -> [p_new (b,n,2) f32, converged (b,n) bool]"""
    rays = _c(rays_img_with_grad, np.float32)
    pts = _c(pts_3d_norm, np.float32)
    p0 = _c(p_init, np.float32)
    b, h, w, c = rays.shape
    assert c == 9
    n = p0.shape[1]
    p_new = np.zeros((b, n, 2), np.float32)
    conv = np.zeros((b, n), np.uint8)
    lib().orc_iter_proj(_p(rays), _p(pts), _p(p0), _p(p_new), _p(conv), b, h, w, n, int(max_iter),
                        C.c_float(lambda_init), C.c_float(cost_thresh))
    return [p_new, conv.astype(bool)]


def refine_matches(D11, D21, p1, radius, dilation_max):
    """reference: refine_matches (gn.cpp:101-114) -> [p1_new (b,n,2) int64].  D11/D21 float16 or float32."""
    D11 = np.ascontiguousarray(D11)
    D21 = np.ascontiguousarray(D21)
    p1 = _c(p1, np.int64)
    b, h, w, fdim = D11.shape
    n = p1.shape[1]
    out = np.zeros((b, n, 2), np.int64)
    if D11.dtype == np.float16:
        assert D21.dtype == np.float16
        a, bb = D11.view(np.uint16), D21.view(np.uint16)
        lib().orc_refine_matches_f16(_p(a), _p(bb), _p(p1), _p(out), b, h, w, n, fdim, int(radius), int(dilation_max))
    else:
        a, bb = _c(D11, np.float32), _c(D21, np.float32)
        lib().orc_refine_matches_f32(_p(a), _p(bb), _p(p1), _p(out), b, h, w, n, fdim, int(radius), int(dilation_max))
    return [out]


def _params(kind, **kw):
    if kind == KIND_POINTS:
        return np.array([kw["sigma_point"]], np.float32)
    if kind == KIND_RAYS:
        return np.array([kw["sigma_ray"], kw["sigma_dist"]], np.float32)
    K = np.asarray(kw["K"], np.float32)
    return np.array([kw["sigma_pixel"], kw["sigma_depth"], K[0, 0], K[1, 1], K[0, 2], K[1, 2], kw["height"],
                     kw["width"], kw["pixel_border"], kw["z_eps"]], np.float32)


def index_prep(ii, jj):
    """get_unique_kf_idx + create_inds(pin=0) (gn_kernels.cu:161-170)"""
    ii = _c(ii, np.int64)
    jj = _c(jj, np.int64)
    E = ii.shape[0]
    uniq = np.zeros(2 * E + 1, np.int64)
    ie = np.zeros(E, np.int64)
    je = np.zeros(E, np.int64)
    nu = lib().orc_index_prep(_p(ii), _p(jj), E, _p(uniq), _p(ie), _p(je))
    return uniq[:nu].copy(), ie, je


def gn_edge_blocks(kind, Twc, Xs, Cs, ii_edge, jj_edge, idx_ii2jj, valid_match, Q, C_thresh, Q_thresh, robust=0,
                   accum=0, **kw):
    """One launch of point_align / ray_align / calib_proj: returns Hs (4,E,7,7), gs (2,E,7)."""
    Twc = _c(Twc, np.float32)
    Xs = _c(Xs, np.float32)
    Cs = _c(Cs, np.float32)
    ie = _c(ii_edge, np.int64)
    je = _c(jj_edge, np.int64)
    idx = _c(idx_ii2jj, np.int64)
    vm = _c(valid_match, np.uint8)
    Q = _c(Q, np.float32)
    num_kf, npts = Xs.shape[0], Xs.shape[1]
    E = ie.shape[0]
    par = _params(kind, **kw)
    Hs = np.zeros((4, E, 7, 7), np.float32)
    gs = np.zeros((2, E, 7), np.float32)
    lib().orc_gn_edge_blocks(kind, _p(par), int(robust), int(accum), _p(Twc), _p(Xs), _p(Cs), _p(ie), _p(je),
                             _p(idx), _p(vm), _p(Q), num_kf, npts, E, C.c_float(C_thresh), C.c_float(Q_thresh),
                             _p(Hs), _p(gs))
    return Hs, gs


def gn_solve(Hs, gs, ii_opt, jj_opt, n_free, return_system=False):
    """SparseBlock.update_lhs/update_rhs/solve, dx = -solve (gn_kernels.cu:57-159,1199-1209)."""
    Hs = _c(Hs, np.float32)
    gs = _c(gs, np.float32)
    io = _c(ii_opt, np.int64)
    jo = _c(jj_opt, np.int64)
    E = io.shape[0]
    dx = np.zeros((n_free, 7), np.float32)
    dim = 7 * n_free
    H = np.zeros((dim, dim), np.float64) if return_system else None
    b = np.zeros((dim,), np.float64) if return_system else None
    fail = lib().orc_gn_solve(_p(Hs), _p(gs), _p(io), _p(jo), E, n_free, _p(dx), _p(H) if return_system else None,
                              _p(b) if return_system else None)
    if return_system:
        return dx, fail, H, b
    return dx, fail


def pose_retr(Twc, dx, num_fix=1):
    Twc = _c(Twc, np.float32).copy()
    dx = _c(dx, np.float32)
    lib().orc_pose_retr(_p(Twc), _p(dx), Twc.shape[0], int(num_fix))
    return Twc


def _gauss_newton(kind, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, C_thresh, Q_thresh, max_iter, delta_thresh,
                  robust=0, accum=0, **kw):
    """Mutates the numpy array `Twc` in place like the reference mutates its tensor. Returns ([dx], iters)."""
    assert Twc.dtype == np.float32 and Twc.flags["C_CONTIGUOUS"]
    Xs = _c(Xs, np.float32)
    Cs = _c(Cs, np.float32)
    ii = _c(ii, np.int64)
    jj = _c(jj, np.int64)
    idx = _c(idx_ii2jj, np.int64)
    vm = _c(valid_match, np.uint8)
    Q = _c(Q, np.float32)
    num_kf, npts = Xs.shape[0], Xs.shape[1]
    E = ii.shape[0]
    par = _params(kind, **kw)
    dx = np.zeros((num_kf - 1, 7), np.float32)
    iters = lib().orc_gauss_newton(kind, _p(par), int(robust), int(accum), _p(Twc), _p(Xs), _p(Cs), _p(ii), _p(jj),
                                   _p(idx), _p(vm), _p(Q), num_kf, npts, E, C.c_float(C_thresh), C.c_float(Q_thresh),
                                   int(max_iter), C.c_float(delta_thresh), _p(dx))
    return [dx], iters


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_point, C_thresh, Q_thresh, max_iter,
                        delta_thresh, **opt):
    """reference: gauss_newton_points (gn.cpp:3-26)"""
    return _gauss_newton(KIND_POINTS, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, C_thresh, Q_thresh, max_iter,
                         delta_thresh, sigma_point=sigma_point, **opt)


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray, sigma_dist, C_thresh, Q_thresh,
                      max_iter, delta_thresh, **opt):
    """reference: gauss_newton_rays (gn.cpp:28-52)"""
    return _gauss_newton(KIND_RAYS, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, C_thresh, Q_thresh, max_iter,
                         delta_thresh, sigma_ray=sigma_ray, sigma_dist=sigma_dist, **opt)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, height, width, pixel_border, z_eps,
                       sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter, delta_thresh, **opt):
    """reference: gauss_newton_calib (gn.cpp:54-82)"""
    return _gauss_newton(KIND_CALIB, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, C_thresh, Q_thresh, max_iter,
                         delta_thresh, K=K, height=height, width=width, pixel_border=pixel_border, z_eps=z_eps,
                         sigma_pixel=sigma_pixel, sigma_depth=sigma_depth, **opt)
