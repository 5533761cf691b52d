"""refine_matches timing at 512x384 on the noisy and the noise-free scene, b = 1 and 8: per-thread gather kernel (mode 0)
against the TMA tile-staged kernel (mode 1); checks that both give the same indices."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import _lib, backends as be, synthetic as syn
dev = "cuda:0"; mt = syn.MATCHING; out = {}
lib = _lib.load()
for noise in (0.005, 0.0):
    for b in (1, 8):
        m = syn.make_match_problem(b, 384, 512, seed=1235, device=dev, point_noise=noise)
        p, c = be.iter_proj(m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"], mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
        D11, D21, p1 = m["D11"].half(), m["D21"].view(b, 384 * 512, -1).half(), p.long()
        fn = lambda: be.refine_matches(D11, D21, p1, mt["radius"], mt["dilation_max"])
        res = {}
        for mode in (0, 1):
            lib.m3r_debug_set_refine_mode(mode)
            for _ in range(3): (q,) = fn()
            torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); e0.record()
            for _ in range(10): fn()
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            res["tile" if mode else "gather"] = dict(ms=round(ms, 4), mpix_s=round(b * 384 * 512 / ms / 1e3, 1))
            if mode: res["equal"] = bool(torch.equal(q, q0))
            else: q0 = q
        out[f"noise{noise}_b{b}"] = res
lib.m3r_debug_set_refine_mode(-1)
print(json.dumps(out, indent=1))
