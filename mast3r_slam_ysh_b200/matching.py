"""Device-side mirror of mast3r_slam/matching.py.

`match(X11, X21, D11, D21, idx_1_to_2_init=None)` has the reference's signature and return values
(matching.py:8-10: `idx_1_to_2` (b, h*w) int64, `valid_match2` (b, h*w, 1) bool) but runs the whole of
`match_iterative_proj` (matching.py:52-91) in four kernels through the C ABI (`m3r_match`, csrc/match_fused.cu):
ray image + Scharr gradients, query normalisation, iter_proj, occlusion test, refine_matches, pixel_to_lin.

`match_unfused` is the reference's own sequence of torch ops around the two drop-in entry points
(`mast3r_slam_backends.iter_proj` / `.refine_matches`): the behaviour a user gets by only swapping the extension.
It is what the fused path is tested and timed against.  No CPU fallback: CPU tensors raise.
"""
import ctypes as C

import torch

from . import _lib
from . import backends as be
from .synthetic import MATCHING, prep_for_iter_proj


def _cfg(cfg):
    c = dict(MATCHING)
    if cfg:
        c.update(cfg)
    return c


def match_unfused(X11, X21, D11, D21, idx_1_to_2_init=None, cfg=None):
    """matching.py:52-91 verbatim in torch, calling the drop-in iter_proj / refine_matches."""
    c = _cfg(cfg)
    b, h, w = X21.shape[:3]
    rays, pts, p_init = prep_for_iter_proj(X11, X21, idx_1_to_2_init)
    p1, valid_proj2 = be.iter_proj(rays, pts, p_init, c["max_iter"], c["lambda_init"], c["convergence_thresh"])
    p1 = p1.long()
    batch_inds = torch.arange(b, device=X11.device)[:, None].repeat(1, h * w)
    dists2 = torch.linalg.norm(X11[batch_inds, p1[..., 1], p1[..., 0], :].reshape(b, h, w, 3) - X21, dim=-1)
    valid_proj2 = valid_proj2 & (dists2 < c["dist_thresh"]).view(b, -1)
    if c["radius"] > 0:
        (p1,) = be.refine_matches(D11.half(), D21.view(b, h * w, -1).half(), p1, c["radius"], c["dilation_max"])
    idx_1_to_2 = p1[..., 0] + (w * p1[..., 1])
    return idx_1_to_2, valid_proj2.unsqueeze(-1)


def match(X11, X21, D11, D21, idx_1_to_2_init=None, cfg=None):
    """Fused device path; same arguments and results as mast3r_slam.matching.match (matching.py:8-10)."""
    c = _cfg(cfg)
    for name, t in (("X11", X11), ("X21", X21)):
        if not t.is_cuda:
            raise RuntimeError(f"match: {name} must be a CUDA tensor (the B200 backend has no CPU fallback)")
        if not t.is_contiguous():
            raise RuntimeError(f"match: {name} must be contiguous")
        if t.dtype != torch.float32:
            raise RuntimeError(f"match: {name} must be float32")
    b, h, w, _ = X11.shape
    n = h * w
    refine = c["radius"] > 0
    if refine:
        D11h = D11.float().contiguous()  # fp32 in: the .half() of matching.py:80-81 happens on the device
        D21h = D21.reshape(b, n, -1).float().contiguous()
        fdim = D11h.shape[-1]
    else:
        D11h = D21h = None
        fdim = 0
    if idx_1_to_2_init is not None:
        idx_1_to_2_init = idx_1_to_2_init.to(torch.int64).contiguous()
    idx = torch.empty((b, n), dtype=torch.int64, device=X11.device)
    valid = torch.empty((b, n), dtype=torch.bool, device=X11.device)
    p = lambda t: C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)
    lib = _lib.load()
    with torch.cuda.device(X11.device):
        rc = lib.m3r_match(p(X11), p(X21), p(D11h), p(D21h), p(idx_1_to_2_init), p(idx), p(valid), b, h, w, fdim,
                           int(c["max_iter"]), float(c["lambda_init"]), float(c["convergence_thresh"]),
                           float(c["dist_thresh"]), int(c["radius"]), int(c["dilation_max"]),
                           C.c_void_p(torch.cuda.current_stream(X11.device).cuda_stream))
    _lib.check(rc, "match")
    return idx, valid.unsqueeze(-1)


def ray_image(X11):
    """Device version of the ray image of prep_for_iter_proj (matching.py:29-36): (b,h,w,3) f32 -> (b,h,w,9) f32."""
    if not X11.is_cuda or X11.dtype != torch.float32 or not X11.is_contiguous():
        raise RuntimeError("ray_image: X11 must be a contiguous float32 CUDA tensor")
    b, h, w, _ = X11.shape
    out = torch.empty((b, h, w, 9), dtype=torch.float32, device=X11.device)
    with torch.cuda.device(X11.device):
        rc = _lib.load().m3r_ray_image(C.c_void_p(X11.data_ptr()), C.c_void_p(out.data_ptr()), b, h, w,
                                       C.c_void_p(torch.cuda.current_stream(X11.device).cuda_stream))
    _lib.check(rc, "ray_image")
    return out
