"""torchrun --nproc-per-node N tools/check_sharded.py : edge-sharded GN vs the single-GPU solve of the same graph, with the
in-kernel peer all-reduce and with the NCCL all-reduce of [H|b]; also checks that pose replicas are bit-identical."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from mast3r_slam_ysh_b200 import backends as be, distributed as md, synthetic as syn
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
lo = syn.LOCAL_OPT
out = {"world": world}
ok = True
# (keyframes, h, w, graph): n = 77 (small single-CTA solver, packed system), n = 273 and n = 553 (envelope solver, row-major
# system; the plan is built from ALL ranks' edges), n = 413 with loop closures everywhere (dense multi-CTA solver)
for nkf, h, w, span in ((12, 96, 128, 3), (40, 48, 64, 3), (80, 24, 32, 8), (60, 24, 32, 0)):
    edges = syn.ring_edges(nkf, span) if span > 0 else syn.loop_closure_edges(nkf, n_loop=3, seed=2)
    ng = len(edges[0])
    g = syn.make_gn_problem(nkf, h, w, edges, seed=31, device=dev, yaw_step_deg=8.0)
    P = (lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], 6, lo["delta_norm"])
    T1 = g["Twc"].clone()
    (dx1,) = be.gauss_newton_rays(T1, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], *P)
    a, b = md.partition_edges(ng, world)[rank]
    sel = md.directed_slice(ng, a, b).to(dev)
    res = {}
    for mode in ("peer", "nccl"):
        T2 = g["Twc"].clone()
        (dx2,), info = md.gauss_newton_rays_sharded(T2, g["Xs"], g["Cs"], g["ii"][sel], g["jj"][sel],
                                                    g["idx_ii2jj"][sel].contiguous(), g["valid_match"][sel].contiguous(),
                                                    g["Q"][sel].contiguous(), g["ii"], g["jj"], *P, exchange=mode)
        torch.cuda.synchronize()
        rel = float((T2 - T1).norm() / T1.norm())
        gath = [torch.empty_like(T2) for _ in range(world)]
        dist.all_gather(gath, T2)
        same = all(torch.equal(gath[0], x) for x in gath)
        res[mode] = {"pose_rel_vs_single_gpu": rel, "replicas_bit_identical": same, "info": info.tolist(),
                     "gt_err": float((T2 - g["Twc_gt"]).abs().max())}
        ok = ok and rel < (1e-5 if span == 3 else 1e-4) and same and (info.tolist()[3] & ~(15 << 24)) == 0
        res[mode + "_T"] = T2
    res["peer_equals_nccl_bits"] = bool(torch.equal(res.pop("peer_T"), res.pop("nccl_T")))
    out[f"nkf{nkf}"] = res
if rank == 0:
    print(json.dumps(out))
    assert ok
dist.destroy_process_group()
