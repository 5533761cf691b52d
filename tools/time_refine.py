"""refine_matches timing on a noisy and a noise-free scene (A/B: M3R_REFINE_AOS=1 uses the (h,w,fdim) layout directly)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import backends as be, synthetic as syn
dev = "cuda:0"; mt = syn.MATCHING; out = {"aos": bool(os.environ.get("M3R_REFINE_AOS"))}
for noise in (0.005, 0.0):
    m = syn.make_match_problem(1, 384, 512, seed=1235, device=dev, point_noise=noise)
    p, c = be.iter_proj(m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"], mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
    D11, D21, p1 = m["D11"].half(), m["D21"].view(1, 384 * 512, -1).half(), p.long()
    fn = lambda: be.refine_matches(D11, D21, p1, mt["radius"], mt["dilation_max"])
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); e0.record()
    for _ in range(20): fn()
    e1.record(); torch.cuda.synchronize()
    out[f"noise_{noise}"] = dict(ms=e0.elapsed_time(e1) / 20, conv_frac=float(c.float().mean()))
print(json.dumps(out))
