"""Envelope (skyline) Cholesky vs the dense multi-CTA Cholesky on the same graph: agreement of dx / poses after one GN
iteration and the time of the FACTOR|BACKSOLVE launch.  The solve time does not depend on the pixel count, so the
scene is tiny.  usage: time_solver.py NKF SPAN [wrap=1] [loops=0]   (SPAN 0: loop_closure_edges).  dev tool."""
import os, sys, json, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import _lib, backends as be, distributed as md, synthetic as syn

nkf = int(sys.argv[1]) if len(sys.argv) > 1 else 256
span = int(sys.argv[2]) if len(sys.argv) > 2 else 8
wrap = int(sys.argv[3]) if len(sys.argv) > 3 else 1
loops = int(sys.argv[4]) if len(sys.argv) > 4 else 0
dev = torch.device("cuda:0"); lo = syn.LOCAL_OPT
if span > 0:
    ii, jj = syn.ring_edges(nkf, span, wrap=bool(wrap))
    g0 = torch.Generator().manual_seed(5)
    for _ in range(loops):  # a few long-range edges on top of the band
        a, b = torch.randint(0, nkf, (2,), generator=g0).tolist()
        if a != b:
            ii.append(a); jj.append(b)
    edges = (ii, jj)
else:
    edges = syn.loop_closure_edges(nkf)
g = syn.make_gn_problem(nkf, 24, 32, edges, seed=1236, device=dev, yaw_step_deg=360.0 / 32)
lib = _lib.load()
P = _lib.GnParams(_lib.M3R_GN_RAYS, 0, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], None, 0, 0, 0, 0.0, 10, -1.0)
res = {"nkf": nkf, "span": span, "wrap": wrap, "loops": loops, "n": 7 * (nkf - 1)}
out = {}
for mode in (1, 0):
    lib.m3r_debug_set_solver(mode)
    T = g["Twc"].clone()
    s = md.ShardedGaussNewton(P, T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], g["all_ii"], g["all_jj"])
    acc = lambda: lib.m3r_gn_accumulate(C.byref(P), C.c_void_p(T.data_ptr()), s.num_kf, s.num_pts, s.E, C.c_void_p(s.ws_ptr), s._stream())
    sol = lambda: lib.m3r_gn_solve_retract(C.byref(P), C.c_void_p(T.data_ptr()), s.num_kf, s.num_pts, s.E, C.c_void_p(s.ws_ptr), s._stream())
    acc(); sol(); torch.cuda.synchronize()
    out[mode] = T.clone()
    res["lanes_mode%d" % mode] = lib.m3r_debug_solver_lanes(C.c_void_p(s.ws_ptr))
    for _ in range(3): acc(); sol()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): acc()
    b.record(); torch.cuda.synchronize(); t_acc = a.elapsed_time(b) / 20
    a.record()
    for _ in range(20): acc(); sol()
    b.record(); torch.cuda.synchronize(); t_both = a.elapsed_time(b) / 20
    name = "dense" if mode == 1 else "envelope"
    res[name + "_solve_ms"] = t_both - t_acc
    res[name + "_acc_ms"] = t_acc
    if mode == 0 and os.environ.get("M3R_TL"):
        cap = 4096
        buf = torch.zeros(2 * cap, dtype=torch.int64, device=dev)
        acc(); torch.cuda.synchronize()
        lib.m3r_debug_set_timeline(C.c_void_p(buf.data_ptr()), cap)
        sol(); torch.cuda.synchronize()
        lib.m3r_debug_set_timeline(None, 0)
        bb = buf.cpu().tolist(); k = bb[0]
        marks = [(bb[2 * (i + 1)], bb[2 * (i + 1) + 1]) for i in range(k)]
        marks.sort(key=lambda x: x[1])
        res["timeline_us"] = [[marks[i][0], round((marks[i][1] - marks[i - 1][1]) / 1e3, 2)] for i in range(1, len(marks))]
    del s
d = (out[0] - out[1]).double()
res["pose_rel_diff"] = float(d.norm() / out[1].double().norm())
res["pose_moved"] = float((out[1] - g["Twc"]).double().norm())
res["finite"] = bool(torch.isfinite(out[0]).all())
lib.m3r_debug_set_solver(0)
# full drop-in call in both modes, 3 iterations
outs = {}
for mode in (1, 0):
    lib.m3r_debug_set_solver(mode)
    T = g["Twc"].clone()
    (dx,), info = be.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"],
                                       lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], 3, 1e-8, return_info=True)
    outs[mode] = (T.clone(), dx.clone(), info.cpu().tolist())
res["dropin_pose_rel_diff"] = float((outs[0][0] - outs[1][0]).double().norm() / outs[1][0].double().norm())
res["dropin_dx_rel_diff"] = float((outs[0][1] - outs[1][1]).double().norm() / max(outs[1][1].double().norm(), 1e-30))
res["info_dense"], res["info_envelope"] = outs[1][2], outs[0][2]
lib.m3r_debug_set_solver(0)
print(json.dumps(res))
