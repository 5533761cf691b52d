"""iter_proj timing at 512x384 for b = 1, 8, 64 on the noisy and the noise-free synthetic scene (CUDA events): the gather
kernel (mode 0) against the tile-staged kernel (mode 1); with a -DM3R_IP_STATS library also the share of corner fetches the
window served.  M3R_ITER_PROJ_WARM selects the iteration after which the window is placed."""
import os, sys, json, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import _lib, backends as be, synthetic as syn
dev = "cuda:0"; mt = syn.MATCHING; out = {"warm": os.environ.get("M3R_ITER_PROJ_WARM", "1")}
lib = _lib.load()
stats = torch.zeros(2, dtype=torch.int64, device=dev)
lib.m3r_debug_set_iter_proj_stats(C.c_void_p(stats.data_ptr()))
for noise, warm_start in ((0.005, False), (0.0, False), (0.005, True)):
    m8 = syn.make_match_problem(8, 384, 512, seed=1235, device=dev, point_noise=noise, warm_start=warm_start)
    for b in (1, 8, 64):
        rep = lambda t: t.repeat((b + 7) // 8, *([1] * (t.dim() - 1)))[:b].contiguous()
        rays, pts, p0 = rep(m8["rays_img_with_grad"]), rep(m8["pts_3d_norm"]), rep(m8["p_init"])
        fn = lambda: be.iter_proj(rays, pts, p0, mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
        res = {}
        for mode in (0, 1):
            lib.m3r_debug_set_iter_proj_mode(mode)
            for _ in range(3): p, c = fn()
            torch.cuda.synchronize(); stats.zero_(); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); e0.record()
            n = 20 if b <= 8 else 5
            for _ in range(n): fn()
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / n
            res["tile" if mode else "gather"] = dict(ms=round(ms, 4), mpix_s=round(b * 384 * 512 / ms / 1e3, 1))
            if mode:
                st = stats.tolist()
                if sum(st): res["window_hit_share"] = round(st[0] / sum(st), 4)
                res["equal"] = bool(torch.equal(p, pg) and torch.equal(c, cg))
            else:
                pg, cg = p, c
        res["conv"] = round(float(c.float().mean()), 3)
        out[f"noise{noise}_{'warm' if warm_start else 'identity'}_b{b}"] = res
lib.m3r_debug_set_iter_proj_mode(-1)
print(json.dumps(out, indent=1))
