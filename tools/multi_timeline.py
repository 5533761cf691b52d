"""Phase timeline of the fused solver launch on rank 0 of a multi-GPU (weak-scaled ring) solve, peer exchange:
torchrun --nproc-per-node N tools/multi_timeline.py.  Prints the globaltimer phases of CTA 0 of one warm iteration -- the
peer all-reduce is the gap between `assemble` and `start-factor` -- and CUDA-event times of begin / hessian / iteration."""
import os, sys, json, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from mast3r_slam_ysh_b200 import _lib, distributed as md, synthetic as syn
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1: dist.init_process_group("nccl", device_id=dev)
lo = syn.LOCAL_OPT; nkf = 32 * world
edges = syn.ring_edges(nkf, 8); n_graph = len(edges[0])
lo_e, hi_e = md.partition_edges(n_graph, world)[rank]
sel = md.directed_slice(n_graph, lo_e, hi_e)
g = syn.make_gn_problem(nkf, 384, 512, edges, seed=1236, device=dev, directed_subset=sel if world > 1 else None, yaw_step_deg=360.0 / 32)
P = _lib.GnParams(_lib.M3R_GN_RAYS, 0, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], None, 0, 0, 0, 0.0, 10, -1.0)
T = g["Twc"].clone()
mk = lambda: md.ShardedGaussNewton(P, T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], g["all_ii"], g["all_jj"])
def ev(): return torch.cuda.Event(enable_timing=True)
def timed(fn, reps):
    if world > 1: dist.barrier()
    torch.cuda.synchronize(); a, b = ev(), ev(); a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / reps
s = mk(); tmp = mk(); del tmp
res = {"world": world, "nkf": nkf, "exchange": s.exchange}
res["begin_ms"] = timed(lambda: mk(), 3)
for _ in range(3): s.iterate()
res["hessian_ms"] = timed(s.hessian_pass, 10)
T.copy_(g["Twc"]); res["iter_ms"] = timed(s.iterate, 10)
lib = s.lib; cap = 4096
buf = torch.zeros(2 * cap, dtype=torch.int64, device=dev)
if world > 1: dist.barrier()
torch.cuda.synchronize()
if rank == 0: lib.m3r_debug_set_timeline(C.c_void_p(buf.data_ptr()), cap)
s.iterate(); torch.cuda.synchronize()
lib.m3r_debug_set_timeline(None, 0)
if rank == 0:
    b = buf.cpu().tolist(); k = b[0]
    marks = [(b[2 * (i + 1)], b[2 * (i + 1) + 1]) for i in range(k)]
    names = {2: "finalize+zero", 3: "barrier", 4: "assemble", 5: "peer all-reduce (+ its two barriers)", 8: "envelope_factor", 6: "finish",
             8100: "nd_interiors(wait)", 8200: "nd_reduce", 8300: "nd_separator_factor", 8400: "nd_separator_backsolve", 8500: "nd_interiors_backsolve(wait)", 7: "small_solve"}
    prev = None; ph = {}
    for tag, t in marks:
        if tag == 1: prev = t; continue
        if 7000 <= tag < 8000 or tag >= 10000: continue
        ph[names.get(tag, str(tag))] = round((t - prev) / 1e3, 2); prev = t
    res["phases_us_rank0"] = ph
    res["segments"] = int(lib.m3r_debug_solver_segments(C.c_void_p(s.ws_ptr)))
    print(json.dumps(res))
if world > 1: dist.destroy_process_group()
