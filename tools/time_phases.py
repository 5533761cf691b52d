"""Warm per-phase timings of the GN loop (CUDA events): begin (pre-pass), hessian, accumulate, solve_retract."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import _lib, distributed as md, synthetic as syn
nkf = int(sys.argv[1]) if len(sys.argv) > 1 else 32
span = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda:0"); lo = syn.LOCAL_OPT
edges = syn.ring_edges(nkf, span) if span > 0 else syn.loop_closure_edges(nkf)
g = syn.make_gn_problem(nkf, 384, 512, edges, seed=1236, device=dev)
P = _lib.GnParams(_lib.M3R_GN_RAYS, 0, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], None, 0, 0, 0, 0.0, 10, -1.0)
def ev(): return torch.cuda.Event(enable_timing=True)
def timed(fn, reps):
    torch.cuda.synchronize(); a, b = ev(), ev(); a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / reps
T = g["Twc"].clone()
_lib.load().m3r_debug_set_solver(int(os.environ.get("M3R_SOLVER_MODE", "0")))  # 2: plain envelope solver, 1: dense
mk = lambda: md.ShardedGaussNewton(P, T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], g["all_ii"], g["all_jj"])
s = mk(); torch.cuda.synchronize()
res = {"nkf": nkf, "E": int(g["ii"].numel()), "n": 7 * (nkf - 1), "solver_mode": int(os.environ.get("M3R_SOLVER_MODE", "0")),
       "segments": int(s.lib.m3r_debug_solver_segments(__import__("ctypes").c_void_p(s.ws_ptr)))}
tmp = mk(); del tmp; torch.cuda.synchronize()  # leave a cached block of workspace size so begin_ms excludes cudaMalloc
res["begin_ms"] = timed(lambda: mk(), 3)
lib = s.lib; import ctypes as C
acc = lambda: _lib.check(lib.m3r_gn_accumulate(C.byref(P), C.c_void_p(T.data_ptr()), s.num_kf, s.num_pts, s.E, C.c_void_p(s.ws_ptr), s._stream()), "acc")
sol = lambda: _lib.check(lib.m3r_gn_solve_retract(C.byref(P), C.c_void_p(T.data_ptr()), s.num_kf, s.num_pts, s.E, C.c_void_p(s.ws_ptr), s._stream()), "sol")
for _ in range(3): acc(); sol()
res["hessian_ms"] = timed(s.hessian_pass, 10)
res["accumulate_ms"] = timed(acc, 10)
def both(): acc(); sol()
T.copy_(g["Twc"]); res["iter_ms"] = timed(both, 10)
acc(); res["solve_ms"] = res["iter_ms"] - res["accumulate_ms"]
res["records"] = int(lib.m3r_gn_record_count(C.c_void_p(s.ws_ptr), s.num_kf, s.num_pts, s.E))
res["pixel_edges"] = int(g["ii"].numel()) * 384 * 512
print(json.dumps(res))
