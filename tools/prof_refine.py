"""A few refine_matches calls at 512x384 for ncu: python tools/prof_refine.py <b> <noise> <mode 0|1>"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import _lib, backends as be, synthetic as syn
b = int(sys.argv[1]) if len(sys.argv) > 1 else 8
noise = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
mode = int(sys.argv[3]) if len(sys.argv) > 3 else 1
mt = syn.MATCHING
m = syn.make_match_problem(b, 384, 512, seed=1235, device="cuda:0", point_noise=noise)
p, c = be.iter_proj(m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"], mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
D11, D21, p1 = m["D11"].half(), m["D21"].view(b, 384 * 512, -1).half(), p.long()
_lib.load().m3r_debug_set_refine_mode(mode)
for _ in range(3):
    (q,) = be.refine_matches(D11, D21, p1, mt["radius"], mt["dilation_max"])
torch.cuda.synchronize()
print("ok", float((q != p1).any(-1).float().mean()))
