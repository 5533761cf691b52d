import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import numpy as np, torch
from mast3r_slam_ysh_b200 import backends as be, synthetic as syn
import mast3r_slam_backends_ref as ref
mt = syn.MATCHING; dev = "cuda:0"
m = syn.make_match_problem(1, 384, 512, seed=1235, device=dev)
rays, pts, p0 = m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"]
args = (mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
p_r, c_r = ref.iter_proj(rays, pts, p0, *args); torch.cuda.synchronize()
for var in (0,):
    os.environ["M3R_ITER_PROJ_VARIANT"] = str(var)
    p, c = be.iter_proj(rays, pts, p0, *args); torch.cuda.synchronize()
    e = (p - p_r).abs().amax(-1)
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(True), torch.cuda.Event(True); a.record(); be.iter_proj(rays, pts, p0, *args); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    print(var, "max", float(e.max()), "n>1e-3", int((e > 1e-3).sum()), "n>0", int((e > 0).sum()), "conv_mis", int((c != c_r).sum()), "ms", min(ts))
