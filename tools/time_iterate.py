import os, sys, json, ctypes as C
sys.path.insert(0, "/root/repo")
import torch
from mast3r_slam_ysh_b200 import _lib, distributed as md, synthetic as syn
nkf = int(sys.argv[1]); dev = torch.device("cuda:0"); lo = syn.LOCAL_OPT
edges = syn.loop_closure_edges(nkf)
g = syn.make_gn_problem(nkf, 384, 512, edges, seed=1236, device=dev)
P = _lib.GnParams(_lib.M3R_GN_RAYS, 0, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], None, 0, 0, 0, 0.0, 10, -1.0)
T = g["Twc"].clone()
s = md.ShardedGaussNewton(P, T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], g["all_ii"], g["all_jj"])
for _ in range(3): s.iterate()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10): s.iterate()
b.record(); torch.cuda.synchronize()
it = a.elapsed_time(b) / 10
a.record()
for _ in range(10): s.hessian_pass()
b.record(); torch.cuda.synchronize()
h = a.elapsed_time(b) / 10
print(json.dumps({"nkf": nkf, "iterate_ms": it, "hessian_ms": h, "fused_all_ms": it - h}))
