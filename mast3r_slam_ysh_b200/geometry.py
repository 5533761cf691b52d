"""Device-side mirror of the element-wise geometry helpers around the hot path.

`constrain_points_to_ray(img_size, Xs, K)` has the signature of mast3r_slam/geometry.py:37-42 and returns a new tensor
of the same shape; one kernel through the C ABI (`m3r_constrain_points_to_ray`), bit-identical to the reference's
torch ops (get_pixel_coords + backproject).  No CPU fallback.
"""
import torch

from . import _lib


def constrain_points_to_ray(img_size, Xs, K, out=None):
    if not Xs.is_cuda or Xs.dtype != torch.float32:
        raise RuntimeError("constrain_points_to_ray: Xs must be a float32 CUDA tensor (the B200 backend has no CPU fallback)")
    h, w = int(img_size[0]), int(img_size[1])
    if Xs.shape[-1] != 3 or Xs.numel() % (3 * h * w) != 0:
        raise RuntimeError(f"constrain_points_to_ray: Xs {tuple(Xs.shape)} does not hold {h}x{w} pointmaps")
    Xc = Xs.contiguous()
    Kc = K.to(device=Xs.device, dtype=torch.float32).contiguous()
    if out is None:
        out = torch.empty_like(Xc)
    frames = Xc.numel() // (3 * h * w)
    with torch.cuda.device(Xs.device):
        rc = _lib.load().m3r_constrain_points_to_ray(Xc.data_ptr(), Kc.data_ptr(), out.data_ptr(), frames, h, w,
                                                     torch.cuda.current_stream(Xs.device).cuda_stream)
    _lib.check(rc, "constrain_points_to_ray")
    return out.view(Xs.shape)


_MODES = {"weighted_pointmap": 0, "indep_conf": 1, "weighted_spherical": 2}


def update_pointmap(X_canon, C, X_new, C_new, T=None, filtering_mode="weighted_pointmap"):
    """Frame.update_pointmap (mast3r_slam/frame.py:41-105) for an already initialised frame (N > 0), IN PLACE on
    `X_canon` (n,3) and the accumulated confidence `C` (n,1); the caller keeps the counters N / N_updates.  `T`: optional
    Sim3 pose ((1,8) tensor or object with `.data`) applied to `X_new` first (tracker.py:96-97)."""
    if filtering_mode not in _MODES:
        raise NotImplementedError(f"update_pointmap: filtering_mode {filtering_mode!r} has no device kernel "
                                  "(first / recent / best_score are plain copies in the reference)")
    for name, t in (("X_canon", X_canon), ("C", C), ("X_new", X_new), ("C_new", C_new)):
        if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
            raise RuntimeError(f"update_pointmap: {name} must be a contiguous float32 CUDA tensor")
    n = C.numel()
    if X_canon.numel() != 3 * n or X_new.numel() != 3 * n or C_new.numel() != n:
        raise RuntimeError("update_pointmap: inconsistent sizes")
    Td = None
    if T is not None:
        Td = (T.data if not isinstance(T, torch.Tensor) else T).to(torch.float32).contiguous()
        if Td.numel() != 8:
            raise RuntimeError("update_pointmap: T must hold one Sim3 pose")
    with torch.cuda.device(X_canon.device):
        rc = _lib.load().m3r_update_pointmap(X_canon.data_ptr(), C.data_ptr(), X_new.data_ptr(), C_new.data_ptr(),
                                             Td.data_ptr() if Td is not None else None, n, _MODES[filtering_mode],
                                             torch.cuda.current_stream(X_canon.device).cuda_stream)
    _lib.check(rc, "update_pointmap")
