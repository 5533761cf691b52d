"""Generates tests/golden/*.npz on a GPU box from the REFERENCE's own CUDA kernels (oracle/_ref, built by
oracle/build_ref.py from /root/reference with the 4-line torch-2.11 API patch).  Run:

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'   # then copy gpurun_out/golden/*.npz here

Each file stores the exact inputs (small synthetic cases from mast3r_slam_ysh_b200/synthetic.py) and the reference
outputs, so the fixtures pin both the CPU oracle (-m "not gpu") and the sm_100a kernels (-m gpu).
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import numpy as np, torch
import mast3r_slam_backends_ref as ref
from mast3r_slam_ysh_b200 import synthetic as syn

out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
os.makedirs(out_dir, exist_ok=True)
dev = "cuda:0"; lo, mt = syn.LOCAL_OPT, syn.MATCHING
c = lambda t: t.detach().cpu().numpy()

# ---- matching: identity init and warm start, two sizes (incl. a non-multiple-of-block size)
for name, (h, w, b, warm, seed) in {"match_a": (40, 56, 2, False, 11), "match_b": (33, 47, 1, True, 12)}.items():
    m = syn.make_match_problem(b, h, w, seed=seed, yaw_step_deg=8.0, warm_start=warm)
    rays, pts, p0 = m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"]
    p_new, conv = ref.iter_proj(rays.to(dev), pts.to(dev), p0.to(dev), mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
    p1 = p_new.long()
    D11 = m["D11"].half(); D21 = m["D21"].view(b, h * w, -1).half()
    (p1n,) = ref.refine_matches(D11.to(dev), D21.to(dev), p1, mt["radius"], mt["dilation_max"])
    # second refine case: radius 2, dilation 3, and start points pushed to the image border
    p1b = p1.clone(); p1b[:, ::7, 0] = 0; p1b[:, ::11, 1] = h - 1
    (p1nb,) = ref.refine_matches(D11.to(dev), D21.to(dev), p1b, 2, 3)
    torch.cuda.synchronize()
    np.savez_compressed(os.path.join(out_dir, name + ".npz"), rays=c(rays), pts=c(pts), p_init=c(p0), p_new=c(p_new),
                        converged=c(conv), D11=c(D11).view(np.uint16), D21=c(D21).view(np.uint16), p1=c(p1), p1_new=c(p1n),
                        p1_b=c(p1b), p1_new_b=c(p1nb), h=h, w=w)
    print(name, "conv", float(conv.float().mean()), "moved", float((p1n != p1).any(-1).float().mean()))

# ---- Gauss-Newton, three kinds; non-contiguous global keyframe ids
for kind in ("rays", "calib", "points"):
    h, w, nkf = 32, 48, 5
    g = syn.make_gn_problem(nkf, h, w, syn.ring_edges(nkf, 2, wrap=False), seed=21, calib=(kind == "calib"),
                            yaw_step_deg=9.0, kf_ids=[3, 4, 7, 9, 12])
    a = [g[k].to(dev) for k in ("Xs", "Cs")]; rest = [g[k].to(dev) for k in ("ii", "jj", "idx_ii2jj", "valid_match", "Q")]
    def run(T, it):
        if kind == "rays": return ref.gauss_newton_rays(T, *a, *rest, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], it, lo["delta_norm"])
        if kind == "calib": return ref.gauss_newton_calib(T, *a, g["K"].to(dev), *rest, h, w, lo["pixel_border"], lo["depth_eps"], lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"], lo["Q_conf"], it, lo["delta_norm"])
        return ref.gauss_newton_points(T, *a, *rest, lo["sigma_point"], lo["C_conf"], lo["Q_conf"], it, lo["delta_norm"])
    T1 = g["Twc"].to(dev).clone(); (dx1,) = run(T1, 1)
    T5 = g["Twc"].to(dev).clone(); (dx5,) = run(T5, 5)
    torch.cuda.synchronize()
    np.savez_compressed(os.path.join(out_dir, f"gn_{kind}.npz"), Twc=c(g["Twc"]), Xs=c(g["Xs"]), Cs=c(g["Cs"]), K=c(g["K"]),
                        ii=c(g["ii"]), jj=c(g["jj"]), idx=c(g["idx_ii2jj"]).astype(np.int32), valid=c(g["valid_match"]),
                        Q=c(g["Q"]), dx1=c(dx1), T1=c(T1), dx5=c(dx5), T5=c(T5), h=h, w=w)
    print(kind, "dx1", float(dx1.norm()), "dx5", float(dx5.norm()))
# ---- pose retraction KAT incl. the small-angle / small-sigma branches (gn_kernels.cu:350-373)
g = torch.Generator().manual_seed(5)
poses = torch.randn(9, 8, generator=g); poses[:, 3:7] = torch.nn.functional.normalize(poses[:, 3:7], dim=-1); poses[:, 7] = poses[:, 7].abs() + 0.5
dx = 0.1 * torch.randn(8, 7, generator=g)
dx[0, 3:6] = 0; dx[0, 6] = 0; dx[1, 3:6] = 1e-8; dx[2, 6] = 1e-9; dx[3, 3:6] = 0; dx[4] = 0
# no standalone binding for pose_retr in the reference: drive it through a GN call is not possible, so this KAT
# is produced by the reference kernel indirectly -> skipped; retraction is covered by T1/T5 above.
print("done ->", out_dir)
