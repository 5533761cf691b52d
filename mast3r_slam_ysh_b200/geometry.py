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
