"""Device-side mirror of the pose solves of mast3r_slam/tracker.py.

`opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid)` and `opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid,
meas_k, valid_meas_k, K, img_size)` have the argument lists of the `FrameTracker` methods of the same name
(tracker.py:173, :216) and return `(T_WCf, T_CkCf)`.  Poses are accepted either as (1,8)/(8,) float32 tensors in
lietorch's `Sim3.data` layout [t, q_xyzw, s] or as objects with a `.data` tensor (a `lietorch.Sim3`); the result has
the type of the input (`type(T)(data)` for objects).  The whole Gauss-Newton loop -- up to `max_iters` iterations with
the reference's convergence test -- is ONE kernel launch through the C ABI (`m3r_track_rays` / `m3r_track_calib`,
csrc/track.cu); the reference runs ~40 eager torch ops and a `.item()` sync per iteration.

`torch.linalg.cholesky` raising inside `FrameTracker.solve` (tracker.py:168; caught at :89-91 to skip the frame) maps to
a RuntimeError here when `check=True` (one 8-byte device->host read after the solve); with `check=False` nothing
synchronises and the status stays on the device in the returned `info` tensor.  No CPU fallback: CPU tensors raise.
"""
import torch

from . import _lib

# config/base.yaml:16-33 (`tracking:`)
TRACKING = dict(max_iters=50, C_conf=0.0, Q_conf=1.5, rel_error=1e-3, delta_norm=1e-3, huber=1.345, sigma_ray=0.003,
                sigma_dist=10.0, sigma_pixel=1.0, sigma_depth=10.0, pixel_border=-10, depth_eps=1e-6)


def _cfg(cfg):
    c = dict(TRACKING)
    if cfg:
        c.update(cfg)
    return c


def _pose_data(T, name):
    d = T.data if not isinstance(T, torch.Tensor) else T
    if not d.is_cuda:
        raise RuntimeError(f"tracker: {name} must be on a CUDA device (the B200 backend has no CPU fallback)")
    if d.numel() != 8:
        raise RuntimeError(f"tracker: {name} must hold one Sim3 pose (8 floats), got shape {tuple(d.shape)}")
    if d.dtype is torch.float32 and d.is_contiguous():
        return d
    return d.detach().to(torch.float32).contiguous()


def _like(T, data):
    return data if isinstance(T, torch.Tensor) else type(T)(data)


def _flat(t, name, n, dtype, cols=1):
    if not t.is_cuda:
        raise RuntimeError(f"tracker: {name} must be a CUDA tensor (the B200 backend has no CPU fallback)")
    if t.numel() != n * cols:
        raise RuntimeError(f"tracker: {name} has {t.numel()} elements, expected {n * cols}")
    if t.dtype is dtype and t.is_contiguous():  # the usual case: no torch op on the call path
        return t
    return t.detach().to(dtype).contiguous()


class TrackInfo:
    """Device-resident outcome of a solve: `info` int32 [iterations, status], `cost` float64 [last 0.5 b^T b]."""

    def __init__(self, buf):
        self._buf = buf  # float64[10]: [0:4] T_WCf, [4:8] T_CkCf (8 f32 each), [8] cost, [9] info (2 int32)

    @property
    def info(self):
        return self._buf[9:10].view(torch.int32)

    @property
    def cost(self):
        return self._buf[8:9]

    @property
    def iterations(self):
        return int(self.info[0].item())

    @property
    def status(self):
        return int(self.info[1].item())


def _ptr(t):
    return t.data_ptr() if t is not None and t.numel() > 0 else None


def _outputs(dev):
    """One allocation for all results; returns the buffer and the four device addresses the C ABI writes to."""
    buf = torch.empty(10, dtype=torch.float64, device=dev)
    base = buf.data_ptr()
    return buf, base, base + 32, base + 72, base + 64  # T_WCf_out, T_CkCf_out, info, cost


def _call(dev, fn, *args):
    """Run the C entry point with `dev` current, on torch's current stream of that device."""
    if dev.index is None or dev.index == torch.cuda.current_device():
        return fn(*args, torch.cuda.current_stream().cuda_stream)
    with torch.cuda.device(dev):
        return fn(*args, torch.cuda.current_stream(dev).cuda_stream)


def _finish(T_WCf, buf, check, return_info, what):
    ti = TrackInfo(buf)
    if check and ti.status != 0:
        raise RuntimeError(f"{what}: Cholesky failed (H is not positive definite)")
    f = buf[0:8].view(torch.float32)
    res = (_like(T_WCf, f[0:8].view(1, 8)), _like(T_WCf, f[8:16].view(1, 8)))
    return res + (ti,) if return_info else res


def opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg=None, check=True, return_info=False):
    """FrameTracker.opt_pose_ray_dist_sim3 (tracker.py:173-214) on the device."""
    c = _cfg(cfg)
    n = Xf.shape[0] if Xf.dim() > 1 else Xf.numel() // 3
    dev = Xf.device
    Xf_ = _flat(Xf, "Xf", n, torch.float32, 3)
    Xk_ = _flat(Xk, "Xk", n, torch.float32, 3)
    Qk_ = _flat(Qk, "Qk", n, torch.float32)
    valid_ = _flat(valid, "valid", n, torch.bool)
    Tf, Tk = _pose_data(T_WCf, "T_WCf"), _pose_data(T_WCk, "T_WCk")
    buf, o_w, o_r, o_info, o_cost = _outputs(dev)
    rc = _call(dev, _lib.load().m3r_track_rays, _ptr(Xf_), _ptr(Xk_), _ptr(Tf), _ptr(Tk), _ptr(Qk_), _ptr(valid_), n,
               c["sigma_ray"], c["sigma_dist"], c["huber"], int(c["max_iters"]), c["rel_error"], c["delta_norm"], o_w,
               o_r, o_info, o_cost)
    _lib.check(rc, "opt_pose_ray_dist_sim3")
    return _finish(T_WCf, buf, check, return_info, "opt_pose_ray_dist_sim3")


def opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size, cfg=None, check=True,
                        return_info=False):
    """FrameTracker.opt_pose_calib_sim3 (tracker.py:216-266) on the device.  `Xk` is unused by the reference's body as
    well (the keyframe enters through `meas_k`); it is kept for the signature."""
    c = _cfg(cfg)
    n = Xf.shape[0] if Xf.dim() > 1 else Xf.numel() // 3
    dev = Xf.device
    Xf_ = _flat(Xf, "Xf", n, torch.float32, 3)
    meas_ = _flat(meas_k, "meas_k", n, torch.float32, 3)
    Qk_ = _flat(Qk, "Qk", n, torch.float32)
    valid_ = _flat(valid, "valid", n, torch.bool)
    vmeas_ = _flat(valid_meas_k, "valid_meas_k", n, torch.bool)
    K_ = _flat(K, "K", 9, torch.float32)
    Tf, Tk = _pose_data(T_WCf, "T_WCf"), _pose_data(T_WCk, "T_WCk")
    buf, o_w, o_r, o_info, o_cost = _outputs(dev)
    rc = _call(dev, _lib.load().m3r_track_calib, _ptr(Xf_), _ptr(meas_), _ptr(Tf), _ptr(Tk), _ptr(Qk_), _ptr(valid_),
               _ptr(vmeas_), _ptr(K_), n, int(img_size[0]), int(img_size[1]), c["sigma_pixel"], c["sigma_depth"],
               c["pixel_border"], c["depth_eps"], c["huber"], int(c["max_iters"]), c["rel_error"], c["delta_norm"], o_w,
               o_r, o_info, o_cost)
    _lib.check(rc, "opt_pose_calib_sim3")
    return _finish(T_WCf, buf, check, return_info, "opt_pose_calib_sim3")
