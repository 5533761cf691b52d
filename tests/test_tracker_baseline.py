"""The CPU baseline of BASELINE.json configs[0] (pure-torch tracker solve, oracle/tracker_torch.py) converges to the
ground-truth relative pose on a synthetic pair and agrees with the graph solver's formulation on one edge."""
import numpy as np
import torch


def make_pair(h, w, seed=2):
    from mast3r_slam_ysh_b200 import synthetic as syn
    g = syn.make_gn_problem(2, h, w, ([0], [1]), seed=seed, yaw_step_deg=8.0, pose_noise=(0.01, 0.01, 0.005))
    # directed edge 0: i=0 (keyframe k), j=1 (frame f): pixel n of f matches pixel idx[n] of k
    idx = g["idx_ii2jj"][0]
    valid = g["valid_match"][0].float() * (g["Q"][0] > 1.5).float()
    return g, g["Xs"][1], g["Xs"][0][idx], g["Q"][0], valid


def test_tracker_port_converges_to_graph_optimum(orc):
    """Same objective as a one-edge graph: the tracker's fixed point equals the graph solver's fixed point."""
    from oracle import tracker_torch as tt
    g, Xf, Xk, Qk, valid = make_pair(36, 48)
    cfg = dict(tt.TRACKING, rel_error=0.0, delta_norm=1e-7)
    T_new, _, iters = tt.opt_pose_ray_dist_sim3(Xf, Xk, tt.Sim3(g["Twc"][1:2].clone()), tt.Sim3(g["Twc"][0:1].clone()),
                                                Qk, valid, cfg=cfg)
    assert 2 <= iters <= 50
    T = g["Twc"].numpy().copy()
    sel = slice(0, 1)
    orc.gauss_newton_rays(T, g["Xs"].numpy(), g["Cs"].numpy(), g["ii"].numpy()[sel], g["jj"].numpy()[sel],
                          g["idx_ii2jj"].numpy()[sel], g["valid_match"].numpy()[sel], g["Q"].numpy()[sel], 0.003, 10.0,
                          0.0, 1.5, 50, 1e-7)
    assert np.allclose(T[1], T_new.data[0].numpy(), rtol=5e-3, atol=5e-4)


def test_tracker_first_step_matches_graph_oracle(orc):
    """Same residuals/weights: one tracker GN step on T_CkCf equals one graph GN step when keyframe k is the fixed
    pose at identity-like gauge (cross-check of the Jacobian conventions)."""
    from oracle import tracker_torch as tt
    g, Xf, Xk, Qk, valid = make_pair(24, 32, seed=5)
    Twc = g["Twc"].clone()
    Twc[0] = torch.tensor([0, 0, 0, 0, 0, 0, 1, 1.0])  # keyframe at identity: world == keyframe frame
    T_new, _, _ = tt.opt_pose_ray_dist_sim3(Xf, Xk, tt.Sim3(Twc[1:2].clone()), tt.Sim3(Twc[0:1].clone()), Qk, valid, max_iters=1)
    T = Twc.numpy().copy()
    sel = slice(0, 1)  # only the directed edge (i=0 <- j=1)
    orc.gauss_newton_rays(T, g["Xs"].numpy(), g["Cs"].numpy(), g["ii"].numpy()[sel], g["jj"].numpy()[sel],
                          g["idx_ii2jj"].numpy()[sel], g["valid_match"].numpy()[sel], g["Q"].numpy()[sel], 0.003, 10.0,
                          0.0, 1.5, 1, 1e-8)
    assert np.allclose(T[1], T_new.data[0].numpy(), rtol=2e-3, atol=2e-4)


def test_calib_tracker_port_converges_towards_ground_truth():
    """opt_pose_calib_sim3 restatement (tracker.py:216-266): pixel + log-depth residuals pull a perturbed frame pose back
    to the ground truth; meas_k / valid_meas_k built as get_points_poses does (tracker.py:141-150)."""
    from mast3r_slam_ysh_b200 import synthetic as syn
    from oracle import tracker_torch as tt
    t = syn.make_track_problem(48, 64, seed=3, calib=True)
    m, vm = tt.make_meas(t["Xk"], t["img_size"])
    assert torch.equal(m, t["meas_k"]) and torch.equal(vm, t["valid_meas_k"])
    Tw, Tr, iters = tt.opt_pose_calib_sim3(t["Xf"], t["Xk"], tt.Sim3(t["T_WCf"].clone()), tt.Sim3(t["T_WCk"].clone()),
                                          t["Qk"], t["valid"], t["meas_k"], t["valid_meas_k"], t["K"], t["img_size"])
    assert 2 <= iters < 50
    e0 = (t["T_WCf"] - t["T_WCf_gt"]).abs().max().item()
    e1 = (Tw.data - t["T_WCf_gt"]).abs().max().item()
    assert e1 < 0.5 * e0
    # fp64 run of the same restatement agrees with the fp32 one (what the device kernel is tested against)
    f = lambda x: x.double() if x.dtype.is_floating_point else x
    Tw64, _, it64 = tt.opt_pose_calib_sim3(f(t["Xf"]), f(t["Xk"]), tt.Sim3(f(t["T_WCf"]).clone()),
                                           tt.Sim3(f(t["T_WCk"]).clone()), f(t["Qk"]), f(t["valid"]), f(t["meas_k"]),
                                           t["valid_meas_k"], f(t["K"]), t["img_size"])
    assert it64 == iters and np.allclose(Tw64.data.numpy(), Tw.data.numpy(), atol=2e-5)
