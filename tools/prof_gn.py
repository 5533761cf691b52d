"""Profiling driver: ONE gauss_newton_rays solve of BASELINE.json configs[2] (or a smaller graph) + one matching pass.
Usage: python tools/prof_gn.py [nkf] [span] [iters]   (run it plain first, then under ncu; see profiles/README.md)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import backends as be, synthetic as syn
nkf = int(sys.argv[1]) if len(sys.argv) > 1 else 32
span = int(sys.argv[2]) if len(sys.argv) > 2 else 8
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 2
dev = "cuda:0"; lo, mt = syn.LOCAL_OPT, syn.MATCHING
g = syn.make_gn_problem(nkf, 384, 512, syn.ring_edges(nkf, span), seed=1236, device=dev)
args = [g[k] for k in ("Xs", "Cs", "ii", "jj", "idx_ii2jj", "valid_match", "Q")]
for rep in range(2):
    T = g["Twc"].clone()
    (dx,), info = be.gauss_newton_rays(T, *args, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], iters, lo["delta_norm"], return_info=True)
torch.cuda.synchronize()
print("gn info", info.tolist(), "dx", float(dx.norm()))
del g, args
m = syn.make_match_problem(1, 384, 512, seed=1235, device=dev)
for rep in range(2):
    p, c = be.iter_proj(m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"], mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
    (q,) = be.refine_matches(m["D11"].half(), m["D21"].view(1, 384 * 512, -1).half(), p.long(), mt["radius"], mt["dilation_max"])
torch.cuda.synchronize()
print("match conv", float(c.float().mean()))
