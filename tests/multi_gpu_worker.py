"""torchrun --nproc-per-node N tests/multi_gpu_worker.py OUT.json : the multi-GPU paths against the single-GPU solve of
the same problem (driven by tests/test_gpu_multi.py; needs N >= 2 devices).

  * edge-sharded Gauss-Newton, rays / calib / points, with the in-kernel peer all-reduce and with the NCCL all-reduce of
    [H | b]: poses equal the single-GPU solve, replicas are bit-identical, peer == NCCL bit for bit;
  * four graph shapes so that all three linear solvers run (small single-CTA, envelope, dense multi-CTA);
  * pair-sharded matching (iter_proj + refine_matches and the fused match): the gathered result equals the single-GPU
    result of all pairs, also with a ragged split.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from mast3r_slam_ysh_b200 import backends as be, distributed as md, matching as mm, synthetic as syn

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
lo, mt = syn.LOCAL_OPT, syn.MATCHING
out = {"world": world}
ok = True


def run_kind(kind, g, sel, exchange):
    T = g["Twc"].clone()
    a = (g["Xs"], g["Cs"])
    loc = (g["ii"][sel], g["jj"][sel], g["idx_ii2jj"][sel].contiguous(), g["valid_match"][sel].contiguous(),
           g["Q"][sel].contiguous(), g["ii"], g["jj"])
    if kind == "rays":
        (dx,), info = md.gauss_newton_rays_sharded(T, *a, *loc, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"],
                                                   lo["Q_conf"], 6, lo["delta_norm"], exchange=exchange)
    elif kind == "points":
        (dx,), info = md.gauss_newton_points_sharded(T, *a, *loc, lo["sigma_point"], lo["C_conf"], lo["Q_conf"], 6,
                                                     lo["delta_norm"], exchange=exchange)
    else:
        (dx,), info = md.gauss_newton_calib_sharded(T, *a, g["K"], *loc, g["h"], g["w"], lo["pixel_border"],
                                                    lo["depth_eps"], lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"],
                                                    lo["Q_conf"], 6, lo["delta_norm"], exchange=exchange)
    torch.cuda.synchronize()
    return T, info


def run_single(kind, g):
    T = g["Twc"].clone()
    a = (g["Xs"], g["Cs"])
    r = (g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"])
    if kind == "rays":
        be.gauss_newton_rays(T, *a, *r, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], 6, lo["delta_norm"])
    elif kind == "points":
        be.gauss_newton_points(T, *a, *r, lo["sigma_point"], lo["C_conf"], lo["Q_conf"], 6, lo["delta_norm"])
    else:
        be.gauss_newton_calib(T, *a, g["K"], *r, g["h"], g["w"], lo["pixel_border"], lo["depth_eps"], lo["sigma_pixel"],
                              lo["sigma_depth"], lo["C_conf"], lo["Q_conf"], 6, lo["delta_norm"])
    torch.cuda.synchronize()
    return T


# (kind, keyframes, h, w, graph): n = 77 (small single-CTA solver, packed system), n = 273 and n = 553 (envelope solver,
# row-major system; the plan is built from ALL ranks' edges), n = 889 (envelope solver, segmented over four interiors),
# n = 413 with loop closures everywhere (dense multi-CTA solver)
CASES = [("rays", 12, 96, 128, 3), ("rays", 40, 48, 64, 3), ("rays", 80, 24, 32, 8), ("rays", 128, 16, 24, 8), ("rays", 60, 24, 32, 0),
         ("calib", 12, 96, 128, 3), ("points", 12, 96, 128, 3)]
for kind, nkf, h, w, span in CASES:
    edges = syn.ring_edges(nkf, span) if span > 0 else syn.loop_closure_edges(nkf, n_loop=3, seed=2)
    ng = len(edges[0])
    g = syn.make_gn_problem(nkf, h, w, edges, seed=31, device=dev, yaw_step_deg=8.0, calib=(kind == "calib"))
    T1 = run_single(kind, g)
    a, b = md.partition_edges(ng, world)[rank]
    sel = md.directed_slice(ng, a, b).to(dev)
    res = {}
    keep = {}
    for mode in ("peer", "nccl"):
        T2, info = run_kind(kind, g, sel, mode)
        rel = float((T2 - T1).norm() / T1.norm())
        gath = [torch.empty_like(T2) for _ in range(world)]
        dist.all_gather(gath, T2)
        same = all(torch.equal(gath[0], x) for x in gath)
        res[mode] = {"pose_rel_vs_single_gpu": rel, "replicas_bit_identical": same, "info": info.tolist(),
                     "solver_lanes": (info.tolist()[3] >> 24) & 15, "solver_segments": (info.tolist()[3] >> 20) & 15,
                     "gt_err": float((T2 - g["Twc_gt"]).abs().max())}
        ok = ok and rel < (1e-5 if span == 3 else 1e-4) and same and (info.tolist()[3] & ~((15 << 24) | (15 << 20))) == 0
        keep[mode] = T2
    res["peer_equals_nccl_bits"] = bool(torch.equal(keep["peer"], keep["nccl"]))
    ok = ok and res["peer_equals_nccl_bits"]
    out[f"{kind}_nkf{nkf}_span{span}"] = res

# ---- pair-sharded matching: 5 pairs (ragged over 2 or 4 ranks) against all pairs on one GPU
npairs, h, w = 5, 96, 128
m = syn.make_match_problem(npairs, h, w, seed=77, device=dev, warm_start=False)
p_lo, p_hi = md.partition_pairs(npairs, world)[rank]
sl = slice(p_lo, p_hi)
D11h, D21h = m["D11"].half(), m["D21"].view(npairs, h * w, -1).half()
p1_all, conv_all = md.match_pairs_sharded(m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"], D11h, D21h, mt)
p1_g, conv_g = md.match_pairs_sharded(m["rays_img_with_grad"][sl].contiguous(), m["pts_3d_norm"][sl].contiguous(),
                                      m["p_init"][sl].contiguous(), D11h[sl].contiguous(), D21h[sl].contiguous(), mt,
                                      gather=True)
idx_all, val_all = mm.match(m["X11"], m["X21"], m["D11"], m["D21"])
idx_g, val_g = md.match_sharded(m["X11"][sl].contiguous(), m["X21"][sl].contiguous(), m["D11"][sl].contiguous(),
                                m["D21"][sl].contiguous(), gather=True)
torch.cuda.synchronize()
out["match_pairs_sharded"] = {"pairs": npairs, "pairs_this_rank": p_hi - p_lo,
                              "unfused_equal": bool(torch.equal(p1_all, p1_g) and torch.equal(conv_all, conv_g)),
                              "fused_equal": bool(torch.equal(idx_all, idx_g) and torch.equal(val_all, val_g))}
ok = ok and out["match_pairs_sharded"]["unfused_equal"] and out["match_pairs_sharded"]["fused_equal"]
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
out["ok"] = bool(flag.item())
if rank == 0:
    print(json.dumps(out))
    if len(sys.argv) > 1:
        os.makedirs(os.path.dirname(os.path.abspath(sys.argv[1])), exist_ok=True)
        json.dump(out, open(sys.argv[1], "w"), indent=1)
dist.destroy_process_group()
sys.exit(0 if out["ok"] else 1)
