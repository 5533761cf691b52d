"""Device tracker solve (csrc/track.cu through the C ABI) against the pure-torch restatement of the reference's
FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 (oracle/tracker_torch.py; tracker.py:156-266).

Tolerance: pose update within 1e-4 relative of the fp64 run of the restatement (north_star's bound for GN pose updates);
the fp32 restatement itself (what the reference computes) sits at ~1e-5 from that truth."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _oracle(t, calib, dtype, cfg=None):
    from oracle import tracker_torch as tt
    c = dict(tt.TRACKING)
    c.update(cfg or {})
    f = lambda x: x.cpu().to(dtype) if x.dtype.is_floating_point else x.cpu()
    Tf, Tk = tt.Sim3(f(t["T_WCf"]).clone()), tt.Sim3(f(t["T_WCk"]).clone())
    if calib:
        Tw, Tr, it = tt.opt_pose_calib_sim3(f(t["Xf"]), f(t["Xk"]), Tf, Tk, f(t["Qk"]), f(t["valid"]), f(t["meas_k"]),
                                            f(t["valid_meas_k"]), f(t["K"]), t["img_size"], cfg=c)
    else:
        Tw, Tr, it = tt.opt_pose_ray_dist_sim3(f(t["Xf"]), f(t["Xk"]), Tf, Tk, f(t["Qk"]), f(t["valid"]), cfg=c)
    return Tw.data.double(), Tr.data.double(), it


def _device(t, calib, cfg=None, **kw):
    from mast3r_slam_ysh_b200 import tracker as trk
    d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in t.items()}
    if calib:
        return trk.opt_pose_calib_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Qk"], d["valid"], d["meas_k"],
                                       d["valid_meas_k"], d["K"], d["img_size"], cfg=cfg, return_info=True, **kw)
    return trk.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Qk"], d["valid"], cfg=cfg,
                                      return_info=True, **kw)


def _rel(a, b, ref0):
    """|a - b| relative to the size of the update |b - ref0| (both (1,8) double)."""
    return ((a - b).norm() / (b - ref0).norm().clamp_min(1e-12)).item()


@pytest.mark.parametrize("calib", [False, True])
@pytest.mark.parametrize("iters", [1, 3])
def test_fixed_iterations_match_restatement(calib, iters):
    from mast3r_slam_ysh_b200 import synthetic as syn
    t = syn.make_track_problem(96, 128, seed=7, calib=calib)
    cfg = dict(max_iters=iters, rel_error=0.0, delta_norm=0.0)
    Tw64, Tr64, it64 = _oracle(t, calib, torch.float64, cfg)
    Tw32, _, _ = _oracle(t, calib, torch.float32, cfg)
    Tw, Tr, info = _device(t, calib, cfg)
    assert info.iterations == iters == it64 and info.status == 0
    T0 = t["T_WCf"].double()
    assert _rel(Tw.cpu().double(), Tw64, T0) < 1e-4
    assert _rel(Tw.cpu().double(), Tw32, T0) < 1e-4
    Tr0 = Tr64 * 0  # relative pose: compare absolutely against its own size
    assert _rel(Tr.cpu().double(), Tr64, Tr0) < 1e-5


@pytest.mark.parametrize("calib", [False, True])
def test_converges_like_restatement_with_reference_thresholds(calib):
    """config/base.yaml thresholds: same iteration count, same pose, close to the ground truth."""
    from mast3r_slam_ysh_b200 import synthetic as syn
    t = syn.make_track_problem(96, 128, seed=11, calib=calib)
    Tw32, _, it32 = _oracle(t, calib, torch.float32)
    Tw, Tr, info = _device(t, calib)
    assert info.status == 0
    assert info.iterations == it32 and 2 <= it32 < 50
    T0 = t["T_WCf"].double()
    assert _rel(Tw.cpu().double(), Tw32, T0) < 1e-4
    gt = t["T_WCf_gt"].double()
    assert (Tw.cpu().double() - gt).abs().max() < 0.03 and (T0 - gt).abs().max() > 0.01
    assert abs(info.cost.item()) > 0


def test_full_resolution_pair_and_determinism():
    """BASELINE.json configs[0]: one 512x384 pair; two runs are bit-identical (fixed summation order, no atomics)."""
    from mast3r_slam_ysh_b200 import synthetic as syn
    t = syn.make_track_problem(384, 512, seed=1234)
    cfg = dict(max_iters=4, rel_error=0.0, delta_norm=0.0)
    Tw32, _, _ = _oracle(t, False, torch.float32, cfg)
    Tw, Tr, info = _device(t, False, cfg)
    Tw2, Tr2, info2 = _device(t, False, cfg)
    assert info.iterations == 4 and info.status == 0
    assert torch.equal(Tw, Tw2) and torch.equal(Tr, Tr2) and info.cost.item() == info2.cost.item()
    assert _rel(Tw.cpu().double(), Tw32, t["T_WCf"].double()) < 1e-4


@pytest.mark.parametrize("n", [64, 255, 257, 1000])
def test_ragged_sizes(n):
    """Point counts that are not a multiple of the 256-point tile, fewer tiles than CTAs."""
    from mast3r_slam_ysh_b200 import synthetic as syn
    t = syn.make_track_problem(48, 64, seed=5)
    sel = torch.linspace(0, t["valid"].numel() - 1, n).long()
    for k in ("Xf", "Xk", "Qk", "valid"):
        t[k] = t[k][sel].contiguous()
    assert t["valid"].sum() >= 16
    cfg = dict(max_iters=2, rel_error=0.0, delta_norm=0.0)
    Tw64, _, _ = _oracle(t, False, torch.float64, cfg)
    Tw, _, info = _device(t, False, cfg)
    assert info.iterations == 2
    assert _rel(Tw.cpu().double(), Tw64, t["T_WCf"].double()) < 1e-4


def test_empty_and_all_invalid_report_cholesky_failure():
    from mast3r_slam_ysh_b200 import synthetic as syn
    from mast3r_slam_ysh_b200 import tracker as trk
    t = syn.make_track_problem(24, 32, seed=2)
    t["valid"] = torch.zeros_like(t["valid"])
    with pytest.raises(RuntimeError, match="Cholesky failed"):
        _device(t, False)
    Tw, Tr, info = _device(t, False, check=False)  # async form: status on the device, pose untouched
    assert info.status == 1 and info.iterations == 0
    z = torch.zeros((0, 3), device="cuda")
    with pytest.raises(RuntimeError, match="Cholesky failed"):
        trk.opt_pose_ray_dist_sim3(z, z, t["T_WCf"].cuda(), t["T_WCk"].cuda(), torch.zeros((0, 1), device="cuda"),
                                   torch.zeros((0, 1), dtype=torch.bool, device="cuda"))


def test_pose_objects_and_cpu_tensors():
    """lietorch.Sim3-like objects (anything with .data and a constructor from data) come back as the same type."""
    from mast3r_slam_ysh_b200 import synthetic as syn
    from mast3r_slam_ysh_b200 import tracker as trk

    class Pose:
        def __init__(self, data):
            self.data = data

    t = syn.make_track_problem(48, 64, seed=9)
    d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in t.items()}
    Tw, Tr = trk.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], Pose(d["T_WCf"]), Pose(d["T_WCk"]), d["Qk"], d["valid"])
    assert isinstance(Tw, Pose) and isinstance(Tr, Pose) and Tw.data.shape == (1, 8)
    Tw2, _ = trk.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Qk"], d["valid"])
    assert torch.equal(Tw.data, Tw2)
    with pytest.raises(RuntimeError, match="CUDA"):
        trk.opt_pose_ray_dist_sim3(t["Xf"], t["Xk"], t["T_WCf"], t["T_WCk"], t["Qk"], t["valid"])
