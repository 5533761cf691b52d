"""Development probe (GPU box): parity of the sm_100a kernels against the CPU oracle and the reference's own CUDA
build (oracle/_ref), plus quick timings.  Not part of the product path."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import numpy as np, torch
from mast3r_slam_ysh_b200 import backends as be, synthetic as syn
from oracle import oracle as orc
try:
    import mast3r_slam_backends_ref as ref
except Exception as ex:  # noqa
    ref = None; print("no _ref:", ex)
dev = "cuda:0"; lo, mt = syn.LOCAL_OPT, syn.MATCHING
out = {}

def timeit(fn, n=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(True), torch.cuda.Event(True); a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return float(np.median(ts))

def stats(a, b):
    e = np.abs(a - b).max(axis=-1); return dict(max=float(e.max()), frac_le_1e3=float((e <= 1e-3).mean()), n_gt=int((e > 1e-3).sum()))

# ---------------- matching
for (h, w, cpu) in [(96, 128, True), (384, 512, False)]:
    m = syn.make_match_problem(1, h, w, seed=1235, device=dev)
    rays, pts, p0 = m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"]
    args = (mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
    p_new, conv = be.iter_proj(rays, pts, p0, *args)
    r = {"conv_frac": float(conv.float().mean())}
    if cpu:
        p_o, c_o = orc.iter_proj(rays.cpu().numpy(), pts.cpu().numpy(), p0.cpu().numpy(), *args)
        r["vs_oracle"] = stats(p_new.cpu().numpy(), p_o); r["conv_mismatch_oracle"] = int((conv.cpu().numpy() != c_o).sum())
    if ref is not None:
        p_r, c_r = ref.iter_proj(rays, pts, p0, *args); torch.cuda.synchronize()
        r["vs_ref"] = stats(p_new.cpu().numpy(), p_r.cpu().numpy()); r["conv_mismatch_ref"] = int((conv != c_r).sum())
        if cpu: r["oracle_vs_ref"] = stats(p_o, p_r.cpu().numpy())
        r["ref_ms"] = timeit(lambda: ref.iter_proj(rays, pts, p0, *args))
    r["ours_ms"] = timeit(lambda: be.iter_proj(rays, pts, p0, *args))
    p1 = p_new.long(); D11 = m["D11"].half(); D21 = m["D21"].view(1, h * w, -1).half()
    (q,) = be.refine_matches(D11, D21, p1, mt["radius"], mt["dilation_max"])
    r["refine_moved_frac"] = float((q != p1).any(-1).float().mean())
    if cpu:
        (q_o,) = orc.refine_matches(D11.cpu().numpy(), D21.cpu().numpy(), p1.cpu().numpy(), mt["radius"], mt["dilation_max"])
        r["refine_mismatch_oracle"] = int((q.cpu().numpy() != q_o).any(-1).sum())
    if ref is not None:
        (q_r,) = ref.refine_matches(D11, D21, p1, mt["radius"], mt["dilation_max"]); torch.cuda.synchronize()
        r["refine_mismatch_ref"] = int((q != q_r).any(-1).sum())
        r["refine_ref_ms"] = timeit(lambda: ref.refine_matches(D11, D21, p1, mt["radius"], mt["dilation_max"]))
    r["refine_ours_ms"] = timeit(lambda: be.refine_matches(D11, D21, p1, mt["radius"], mt["dilation_max"]))
    out[f"match_{h}x{w}"] = r; print(json.dumps({f"match_{h}x{w}": r}), flush=True)

# ---------------- GN
def gn_case(name, kind, nkf, h, w, edges, cpu, iters=10, **kw):
    g = syn.make_gn_problem(nkf, h, w, edges, seed=1236, device=dev, calib=(kind == "calib"), **kw)
    a = [g[k] for k in ("Xs", "Cs")]; rest = [g[k] for k in ("ii", "jj", "idx_ii2jj", "valid_match", "Q")]
    def run(mod, T, it):
        if kind == "rays": return mod.gauss_newton_rays(T, *a, *rest, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], it, lo["delta_norm"])
        if kind == "calib": return mod.gauss_newton_calib(T, *a, g["K"], *rest, h, w, lo["pixel_border"], lo["depth_eps"], lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"], lo["Q_conf"], it, lo["delta_norm"])
        return mod.gauss_newton_points(T, *a, *rest, lo["sigma_point"], lo["C_conf"], lo["Q_conf"], it, lo["delta_norm"])
    r = {"E": int(g["ii"].numel()), "valid_frac": float(g["valid_match"].float().mean())}
    T1 = g["Twc"].clone(); (dx1,) = run(be, T1, 1); torch.cuda.synchronize()
    Tn = g["Twc"].clone(); (dxn,) = run(be, Tn, iters); torch.cuda.synchronize()
    r["gt_err_init"] = float((g["Twc"] - g["Twc_gt"]).abs().max()); r["gt_err_final"] = float((Tn - g["Twc_gt"]).abs().max())
    if ref is not None:
        R1 = g["Twc"].clone(); (rdx1,) = run(ref, R1, 1); torch.cuda.synchronize()
        Rn = g["Twc"].clone(); (rdxn,) = run(ref, Rn, iters); torch.cuda.synchronize()
        r["dx1_rel_vs_ref"] = float((dx1 - rdx1).norm() / rdx1.norm()); r["T1_rel_vs_ref"] = float((T1 - R1).norm() / R1.norm())
        r["Tn_rel_vs_ref"] = float((Tn - Rn).norm() / Rn.norm())
        r["ref_ms_per_iter"] = timeit(lambda: run(ref, g["Twc"].clone(), iters), n=3, warm=1) / iters
    if cpu:
        an = [x.cpu().numpy() for x in a]; rn = [x.cpu().numpy() for x in rest]
        O1 = g["Twc"].cpu().numpy().copy()
        if kind == "rays": (odx1,), _ = orc.gauss_newton_rays(O1, *an, *rn, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], 1, lo["delta_norm"])
        elif kind == "calib": (odx1,), _ = orc.gauss_newton_calib(O1, *an, g["K"].cpu().numpy(), *rn, h, w, lo["pixel_border"], lo["depth_eps"], lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"], lo["Q_conf"], 1, lo["delta_norm"])
        else: (odx1,), _ = orc.gauss_newton_points(O1, *an, *rn, lo["sigma_point"], lo["C_conf"], lo["Q_conf"], 1, lo["delta_norm"])
        r["dx1_rel_vs_oracle"] = float(np.linalg.norm(dx1.cpu().numpy() - odx1) / np.linalg.norm(odx1))
        if ref is not None: r["oracle_dx1_rel_vs_ref"] = float(np.linalg.norm(rdx1.cpu().numpy() - odx1) / np.linalg.norm(odx1))
    r["ours_ms_per_iter"] = timeit(lambda: run(be, g["Twc"].clone(), iters), n=3, warm=1) / iters
    out[name] = r; print(json.dumps({name: r}), flush=True)

for kind in ("rays", "calib", "points"):
    gn_case(f"gn_{kind}_small", kind, 6, 48, 64, syn.ring_edges(6, 2, wrap=False), True, yaw_step_deg=10.0)
gn_case("gn_rays_10kf_full", "rays", 10, 384, 512, syn.ring_edges(10, 3, wrap=False), False, yaw_step_deg=10.0)
gn_case("gn_calib_10kf_full", "calib", 10, 384, 512, syn.ring_edges(10, 3, wrap=False), False, yaw_step_deg=10.0)
if "--big" in sys.argv:
    gn_case("gn_rays_32kf_full", "rays", 32, 384, 512, syn.ring_edges(32, 8), False)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gpu_check.json"), "w"), indent=1)
