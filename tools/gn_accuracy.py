"""Who is closer to the fp64 truth at full resolution: the reference CUDA (fp32 strided accumulation) or ours?"""
import os, sys, json, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import numpy as np, torch
from mast3r_slam_ysh_b200 import backends as be, synthetic as syn
from oracle import oracle as orc
import mast3r_slam_backends_ref as ref
dev = "cuda:0"; lo = syn.LOCAL_OPT
h, w, nkf = 384, 512, 6
g = syn.make_gn_problem(nkf, h, w, syn.ring_edges(nkf, 2, wrap=False), seed=1236, device=dev, yaw_step_deg=10.0)
a = [g[k] for k in ("Xs", "Cs")]; rest = [g[k] for k in ("ii", "jj", "idx_ii2jj", "valid_match", "Q")]
P = (lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], 1, lo["delta_norm"])
T = g["Twc"].clone(); (dx,) = be.gauss_newton_rays(T, *a, *rest, *P)
R = g["Twc"].clone(); (rdx,) = ref.gauss_newton_rays(R, *a, *rest, *P); torch.cuda.synchronize()
an = [x.cpu().numpy() for x in a]; rn = [x.cpu().numpy() for x in rest]
res = {}
for accum in (0, 1):
    O = g["Twc"].cpu().numpy().copy(); t = time.time()
    (odx,), _ = orc.gauss_newton_rays(O, *an, *rn, *P, accum=accum)
    nrm = np.linalg.norm(odx)
    res[f"accum{accum}"] = dict(sec=time.time() - t, ours_rel=float(np.linalg.norm(dx.cpu().numpy() - odx) / nrm),
                                ref_rel=float(np.linalg.norm(rdx.cpu().numpy() - odx) / nrm))
res["ours_vs_ref"] = float((dx - rdx).norm() / rdx.norm())
# block-level comparison against the fp64-accumulating oracle
uniq, ie, je = orc.index_prep(rn[0], rn[1])
Hs_o, gs_o = orc.gn_edge_blocks(orc.KIND_RAYS, g["Twc"].cpu().numpy(), *an, ie, je, rn[2], rn[3], rn[4], lo["C_conf"], lo["Q_conf"], sigma_ray=lo["sigma_ray"], sigma_dist=lo["sigma_dist"])
Hs, gs = be.gn_edge_blocks("rays", g["Twc"], *a, torch.from_numpy(ie).to(dev), torch.from_numpy(je).to(dev), *rest[2:], lo["C_conf"], lo["Q_conf"], sigma_ray=lo["sigma_ray"], sigma_dist=lo["sigma_dist"])
res["Hs_rel_vs_oracle64"] = float(np.linalg.norm(Hs.cpu().numpy() - Hs_o) / np.linalg.norm(Hs_o))
res["gs_rel_vs_oracle64"] = float(np.linalg.norm(gs.cpu().numpy() - gs_o) / np.linalg.norm(gs_o))
print(json.dumps(res, indent=1))
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "gn_accuracy.json"), "w"), indent=1)
