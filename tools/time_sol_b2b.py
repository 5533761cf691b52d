"""Is the fused solver launch slower when another kernel ran in between (cold instruction cache)?  Times the
FACTOR|BACKSOLVE launch back to back vs alternating with the Hessian/assembly launch.  dev tool."""
import os, sys, json, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import _lib, distributed as md, synthetic as syn
nkf = int(sys.argv[1]) if len(sys.argv) > 1 else 32
span = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda:0"); lo = syn.LOCAL_OPT
edges = syn.ring_edges(nkf, span) if span > 0 else syn.loop_closure_edges(nkf)
g = syn.make_gn_problem(nkf, 384, 512, edges, seed=1236, device=dev)
P = _lib.GnParams(_lib.M3R_GN_RAYS, 0, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], None, 0, 0, 0, 0.0, 10, -1.0)
T = g["Twc"].clone()
s = md.ShardedGaussNewton(P, T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], g["all_ii"], g["all_jj"])
lib = s.lib
acc = lambda: lib.m3r_gn_accumulate(C.byref(P), C.c_void_p(T.data_ptr()), s.num_kf, s.num_pts, s.E, C.c_void_p(s.ws_ptr), s._stream())
sol = lambda: lib.m3r_gn_solve_retract(C.byref(P), C.c_void_p(T.data_ptr()), s.num_kf, s.num_pts, s.E, C.c_void_p(s.ws_ptr), s._stream())
def timed(fn, reps=20):
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / reps
for _ in range(3): acc(); sol()
res = {"nkf": nkf}
res["acc_ms"] = timed(acc)
res["sol_back_to_back_ms"] = timed(sol)
res["acc_plus_sol_ms"] = timed(lambda: (acc(), sol()))
res["sol_after_acc_ms"] = res["acc_plus_sol_ms"] - res["acc_ms"]
print(json.dumps(res))
