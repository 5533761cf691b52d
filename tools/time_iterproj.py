"""iter_proj timing at 512x384 for b = 1, 8, 64 on the noisy and the noise-free synthetic scene (CUDA events)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import backends as be, synthetic as syn
dev = "cuda:0"; mt = syn.MATCHING; out = {}
for noise in (0.005, 0.0):
    m8 = syn.make_match_problem(8, 384, 512, seed=1235, device=dev, point_noise=noise)
    for b in (1, 8, 64):
        rep = lambda t: t.repeat((b + 7) // 8, *([1] * (t.dim() - 1)))[:b].contiguous()
        rays, pts, p0 = rep(m8["rays_img_with_grad"]), rep(m8["pts_3d_norm"]), rep(m8["p_init"])
        fn = lambda: be.iter_proj(rays, pts, p0, mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
        for _ in range(3): p, c = fn()
        torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); e0.record()
        n = 20 if b <= 8 else 5
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        out[f"noise{noise}_b{b}"] = dict(ms=round(ms, 4), mpix_s=round(b * 384 * 512 / ms / 1e3, 1), conv=round(float(c.float().mean()), 3))
print(json.dumps(out, indent=1))
