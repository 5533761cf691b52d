"""A few iter_proj calls at 512x384 for ncu: python tools/prof_iterproj.py <b> <noise> [warm_start]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import backends as be, synthetic as syn
b = int(sys.argv[1]) if len(sys.argv) > 1 else 64
noise = float(sys.argv[2]) if len(sys.argv) > 2 else 0.005
warm = len(sys.argv) > 3 and sys.argv[3] == "1"
mt = syn.MATCHING
m8 = syn.make_match_problem(8, 384, 512, seed=1235, device="cuda:0", point_noise=noise, warm_start=warm)
rep = lambda t: t.repeat((b + 7) // 8, *([1] * (t.dim() - 1)))[:b].contiguous()
rays, pts, p0 = rep(m8["rays_img_with_grad"]), rep(m8["pts_3d_norm"]), rep(m8["p_init"])
for _ in range(3):
    p, c = be.iter_proj(rays, pts, p0, mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
torch.cuda.synchronize()
print("ok", float(c.float().mean()))
