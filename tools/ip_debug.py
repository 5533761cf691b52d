"""One tile-staged iter_proj call on a small image, compared with the gather kernel (run under compute-sanitizer when a
TMA change misbehaves)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import _lib, backends as be, synthetic as syn
lib = _lib.load(); mt = syn.MATCHING
h, w = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (96, 128)
m = syn.make_match_problem(1, h, w, seed=5, device="cuda:0")
a = (m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"], mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
lib.m3r_debug_set_iter_proj_mode(0); pa, ca = be.iter_proj(*a); torch.cuda.synchronize()
lib.m3r_debug_set_iter_proj_mode(1); pb, cb = be.iter_proj(*a); torch.cuda.synchronize()
print("equal", bool(torch.equal(pa, pb) and torch.equal(ca, cb)), "conv", float(ca.float().mean()))
