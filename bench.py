#!/usr/bin/env python
"""bench.py -- GN solve ms/iter (+ iter_proj / refine_matches Mpix/s) at 512x384 on N B200s.

Contract (one JSON line on rank 0):
  metric  gn_solve_ms_per_iter: one "step" = one gauss_newton_rays solve (max_iter = 10 GN iterations: Hessian
          assembly + reduction + fp64 Cholesky + retraction + convergence flag, pre-pass included) of BASELINE.json
          configs[2] (32 keyframes per GPU, 8 edges/kf -> 512 directed edges per GPU, dense 512x384 points), value = step
          time / iterations executed.  N > 1: weak scaling -- 32*N keyframes, 512*N directed edges sharded by edge, the
          fp64 [H | b] summed inside the solver kernel over peer-mapped memory (NCCL all-reduce as fallback).
  e2e     same metric through the reference-facing Python call with pinned HOST buffers: H2D of every input and D2H of
          the poses/dx inside the timed region.
  roofline  the per-record Hessian kernel timed alone with CUDA events.  Unit = one surviving (pixel, directed edge)
          record of the packing pre-pass, 32 B streamed per record and iteration; achieved = records x
          32 B / launch time against MEASURED_PEAKS.json hbm_gbs; `traffic` = DRAM bytes of the same launch from the
          committed ncu capture.  The reference layout's 45 B per pixel-edge is kept as
          `reference_layout_equivalent_gbs` -- it is not a roofline fraction because gated-out pixels are never read
          again after the pre-pass.
  cpu_baseline  the CPU oracle port (oracle/m3r_oracle.c, OpenMP, all host cores) on a bounded sample of the same
          workload, scaled to the full edge count (`extrapolated: true`).
  extra   configs[1] (iter_proj / refine_matches / match at b = 1, 8, 64, 256), configs[3] (gauss_newton_calib),
          configs[4] (fixed 256-keyframe loop-closure graph + 256 pair matches, strong-scaled over the N ranks),
          reference_cuda (the reference's own CUDA kernels, oracle/_ref, on the same GPU and inputs), cpu_tracker
          (configs[0]), tracker, graph_prep, a dense-overlap Hessian point.
  --impl reference  times the CPU port as the reference arm (the reference's GN kernels are CUDA-only; its CPU code
          for this math is the tracker solve, reported as extra.cpu_tracker).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W = 384, 512
N_PIX = H * W
KF_PER_GPU, SPAN, GN_ITERS = 32, 8, 10
RECORD_BYTES = 32               # two float4 per surviving (pixel, edge) record, streamed once per iteration
REF_LAYOUT_BYTES_PER_PIXEL_EDGE = 45  # valid 1 + idx 8 + X_j 12 + X_i 12 + Q 4 + C_i 4 + C_j 4
INPUT_BYTES_PER_PIXEL_EDGE = 13       # idx 8 + valid 1 + Q 4, read once per solve by the pre-pass
ITER_PROJ_BYTES_PER_QUERY = 65  # 36 ray image + 12 target + 8 init + 8 out + 1 flag


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons sampled through NVML every ~2 ms while the timed regions run (the timed GN steps last
    tens of milliseconds, far less than one `nvidia-smi` process start-up, so the query is made in-process)."""
    REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.sm, self.mask, self.max_mhz, self.stop_flag, self.err = index, [], 0, None, False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(t.strip().isdigit() for t in vis.split(",")) else index
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            # the first query of each kind can take longer than the whole timed region: pay for it here
            pynvml.nvmlDeviceGetClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            pynvml.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception as ex:  # noqa
            self.nv, self.err = None, repr(ex)

    def run(self):
        if self.nv is None:
            return
        while not self.stop_flag:
            try:
                self.sm.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                self.mask |= int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
            except Exception as ex:  # noqa
                self.err = repr(ex)
            time.sleep(0.002)

    def summary(self):
        self.stop_flag = True
        if self.is_alive():
            self.join(timeout=1.0)  # a query that was in flight when the timed region ended still counts: it began inside it
        if self.nv is not None and not self.sm:
            # the timed region (a dozen milliseconds) was shorter than NVML's first answer: one query right after it, while
            # the clocks have not had time to drop (flagged below)
            try:
                self.sm.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                self.mask |= int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                self.err = (self.err or "") + "no sample completed inside the timed region; one taken right after it"
            except Exception as ex:  # noqa
                self.err = repr(ex)
        sm = sorted(self.sm)
        out = {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz,
               "reasons": [n for n, bit in self.REASONS.items() if self.mask & bit], "samples": len(sm),
               "source": "NVML, sampled during the timed GN steps and the Hessian-kernel timing loop"}
        if self.err:
            out["error"] = self.err
        return out


def gn_params():
    from mast3r_slam_ysh_b200 import synthetic as syn
    lo = syn.LOCAL_OPT
    return (lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], GN_ITERS, lo["delta_norm"])


# ====================================================================================================== reference arm
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_gn_sample(n_gpus=1, device="cpu"):
    """Times the CPU oracle port on a bounded sample of the workload bench.py --gpus n_gpus names: S of its directed
    edges, ONE GN iteration of edge-block evaluation (the O(E*N) part), scaled to all edges, plus the dense fp64 solve of
    its system.  All host cores, set explicitly (a launcher may export OMP_NUM_THREADS=1)."""
    import numpy as np
    import torch
    from mast3r_slam_ysh_b200 import synthetic as syn
    from oracle import oracle as orc
    lo = syn.LOCAL_OPT
    cores = orc.set_num_threads(host_cores())
    nkf = KF_PER_GPU * n_gpus
    edges = syn.ring_edges(nkf, SPAN)
    E_full = 2 * len(edges[0])
    S = min(E_full, max(8 * cores, 64))  # a few seconds of CPU work per step; the rest of the edges by scaling
    g = syn.make_gn_problem(nkf, H, W, edges, seed=1236, device=device, directed_subset=torch.arange(S),
                            yaw_step_deg=360.0 / KF_PER_GPU)
    a = [g[k].cpu().numpy() for k in ("Twc", "Xs", "Cs")]
    ie, je = g["ii"].cpu().numpy(), g["jj"].cpu().numpy()
    rest = [g[k].cpu().numpy() for k in ("idx_ii2jj", "valid_match", "Q")]

    def one(n_edges):
        t = time.perf_counter()
        orc.gn_edge_blocks(orc.KIND_RAYS, *a, ie[:n_edges], je[:n_edges], rest[0][:n_edges], rest[1][:n_edges],
                           rest[2][:n_edges], lo["C_conf"], lo["Q_conf"], accum=1, sigma_ray=lo["sigma_ray"],
                           sigma_dist=lo["sigma_dist"])
        return time.perf_counter() - t
    one(min(S, cores))  # warm
    t_s = one(S)
    # dense solve part on a random SPD system of the workload's dimension (same code path as the oracle's solve)
    rng = np.random.default_rng(0)
    Hs = np.zeros((4, nkf - 1, 7, 7), np.float32)
    Mb = rng.standard_normal((nkf - 1, 7, 7)).astype(np.float32)
    Hs[0] = Mb @ Mb.transpose(0, 2, 1) + 7 * np.eye(7, dtype=np.float32)
    gs = rng.standard_normal((2, nkf - 1, 7)).astype(np.float32)
    io = np.arange(nkf - 1)
    t = time.perf_counter()
    orc.gn_solve(Hs, gs, io, io - 10 ** 6, nkf - 1)
    t_solve = time.perf_counter() - t
    ms_iter = (t_s * E_full / S + t_solve) * 1e3
    return dict(value=ms_iter, unit="ms/iter", cores=cores, kind="port", extrapolated=True, scale=E_full / S,
                measured_sample_seconds=t_s + t_solve,
                sample=f"{S} of {E_full} directed edges x {N_PIX} px of the {nkf}-keyframe graph, 1 GN iteration of "
                       f"oracle/m3r_oracle.c (fp32 reference-order accumulation, OpenMP x{cores}), edge part scaled "
                       f"x{E_full / S:.0f}, + dense fp64 solve (n = {7 * (nkf - 1)})"), t_s


def cpu_tracker_baseline():
    """BASELINE.md 3b / configs[0]: pure-torch tracker solve (oracle/tracker_torch.py) on one synthetic 512x384 pair on the
    host cores: 1 thread and all cores, median of 20 solves after 3 warm-ups, core count in the report."""
    import torch
    from mast3r_slam_ysh_b200 import synthetic as syn
    from oracle import tracker_torch as tt
    g = syn.make_gn_problem(2, H, W, ([0], [1]), seed=1234, yaw_step_deg=8.0)
    idx = g["idx_ii2jj"][0]
    valid = g["valid_match"][0].float() * (g["Q"][0] > 1.5).float()
    Xf, Xk, Qk = g["Xs"][1], g["Xs"][0][idx], g["Q"][0]
    out = dict(what="tracker.opt_pose_ray_dist_sim3 port (oracle/tracker_torch.py), 1 pair 512x384, torch CPU, "
                    "config/base.yaml thresholds; median of 20 solves after 3 warm-ups",
               host_cores=host_cores())
    keep = torch.get_num_threads()
    for label, nt in (("threads_1", 1), ("threads_all", host_cores())):
        torch.set_num_threads(nt)
        ts, its = [], 0
        for rep in range(23):
            t = time.perf_counter()
            _, _, its = tt.opt_pose_ray_dist_sim3(Xf, Xk, tt.Sim3(g["Twc"][1:2].clone()), tt.Sim3(g["Twc"][0:1].clone()),
                                                  Qk, valid)
            if rep >= 3:
                ts.append(time.perf_counter() - t)
        ts.sort()
        med = ts[len(ts) // 2]
        out[label] = dict(threads=torch.get_num_threads(), solve_ms=med * 1e3, ms_per_iter=med * 1e3 / max(its, 1),
                          iters=its, p10_ms=ts[len(ts) // 10] * 1e3, p90_ms=ts[(9 * len(ts)) // 10] * 1e3)
    torch.set_num_threads(keep)
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    base = None
    for s in range(args.warmup + args.steps):
        base, _ = cpu_gn_sample(args.gpus)
        if s >= args.warmup:
            vals.append(base["value"])
    v = sum(vals) / len(vals)
    base["value"] = v
    line = dict(impl="reference", metric="gn_solve_ms_per_iter", value=v, unit="ms/iter", n_gpus=args.gpus,
                steps=args.steps, warmup=args.warmup, ms_per_step=v * GN_ITERS, higher_is_better=False, scaling="weak",
                vs_baseline=None, dtype="f32 (fp64 solve)", data="synthetic",
                config=workload_config(args.gpus), cpu_baseline=base, extrapolated=True,
                note="CPU port of the reference's CUDA-only GN path (oracle/m3r_oracle.c) on the host cores; every step "
                     "times a bounded edge sample of the named workload and scales the edge part to all edges -- value "
                     "and ms_per_step are extrapolations, not wall-clock of a full solve",
                e2e=dict(value=v, unit="ms/iter", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), flush=True)


def workload_config(n):
    return dict(workload="FactorGraph.solve_GN_rays: %d keyframes, 8 edges/kf -> %d directed edges, dense 512x384 points, "
                         "max_iter 10 (BASELINE.json configs[2]%s)" % (KF_PER_GPU * n, 2 * KF_PER_GPU * SPAN * n,
                                                                      "" if n == 1 else ", weak-scaled at the same 11.25 deg yaw step; N=8 == configs[4] size"),
                keyframes=KF_PER_GPU * n, directed_edges=2 * KF_PER_GPU * SPAN * n, pixels=N_PIX, gn_iters=GN_ITERS,
                parallelism=("edge-sharded x%d, [H|b] fp64 summed %s, replicated Cholesky"
                             % (n, "on one GPU" if n == 1 else
                                ("by NCCL all-reduce" if os.environ.get("M3R_EXCHANGE") == "nccl"
                                 else "inside the solver kernel over peer-mapped memory (NCCL all-reduce as fallback)"))),
                l2_policy="inputs (>1.3 GB per GPU) exceed the 126 MB L2; no explicit flush")


# ====================================================================================================== our arm (GPU)
def _event_time(torch, fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def hessian_point(torch, md, _lib, P, g, kind="rays", extra_params=None, reps=20):
    """The Hessian kernel alone on graph `g`: launch time (CUDA events on the launch stream), surviving records."""
    import ctypes as C
    if kind == "rays":
        params = _lib.GnParams(_lib.M3R_GN_RAYS, _lib.M3R_ROBUST_HUBER, P[0], P[1], P[2], P[3], None, 0, 0, 0, 0.0,
                               GN_ITERS, P[5])
    else:
        params = extra_params
    sh = md.ShardedGaussNewton(params, g["Twc"].clone(), g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                               g["valid_match"], g["Q"], g["all_ii"], g["all_jj"], exchange="nccl")
    ms = _event_time(torch, sh.hessian_pass, reps)
    recs = int(sh.lib.m3r_gn_record_count(C.c_void_p(sh.ws_ptr), sh.num_kf, sh.num_pts, sh.E))
    del sh
    return ms, recs


def run_ours(args):
    import torch
    import torch.distributed as dist
    from mast3r_slam_ysh_b200 import _lib, backends as be, distributed as md, synthetic as syn

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 backend has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    P = gn_params()
    nkf = KF_PER_GPU * world
    edges = syn.ring_edges(nkf, SPAN)
    n_graph = len(edges[0])
    lo_e, hi_e = md.partition_edges(n_graph, world)[rank]
    sel = md.directed_slice(n_graph, lo_e, hi_e)
    # the yaw step between consecutive keyframes stays that of the 32-keyframe circle (11.25 deg) for every N, so an
    # edge sees the same overlap (same share of valid matches) at every scale: per-GPU work is fixed (weak scaling)
    g = syn.make_gn_problem(nkf, H, W, edges, seed=1236, device=dev, directed_subset=sel if world > 1 else None,
                            yaw_step_deg=360.0 / KF_PER_GPU)
    Xs, Cs, ii, jj, idx, vm, Q = (g[k] for k in ("Xs", "Cs", "ii", "jj", "idx_ii2jj", "valid_match", "Q"))
    T0 = g["Twc"].clone()
    T = T0.clone()
    E_local, E_total = int(ii.numel()), 2 * n_graph
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    info_keep = {}

    def step():
        T.copy_(T0)
        if world == 1:
            (dx,), info = be.gauss_newton_rays(T, Xs, Cs, ii, jj, idx, vm, Q, *P, return_info=True)
        else:
            (dx,), info = md.gauss_newton_rays_sharded(T, Xs, Cs, ii, jj, idx, vm, Q, g["all_ii"], g["all_jj"], *P)
        info_keep["info"] = info
        return dx

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = lib.m3r_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    launches = int(lib.m3r_launch_count() - l0)
    total_ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    iters = int(info_keep["info"].cpu()[0])
    ms_per_step = total_ms / args.steps
    value = ms_per_step / max(iters, 1)

    # ---- dominant kernel alone: the per-record Hessian kernel over this rank's edges, CUDA events on the launch stream
    hess_ms, records = hessian_point(torch, md, _lib, P, dict(g, Twc=T0))
    clocks = sampler.summary() if rank == 0 else None
    peak, peak_src = measured_peak()
    algo_bytes = RECORD_BYTES * records
    achieved = algo_bytes / (hess_ms * 1e-3) / 1e9
    pixel_edges = N_PIX * E_local
    traffic = ncu_traffic()
    input_bytes = INPUT_BYTES_PER_PIXEL_EDGE * pixel_edges + 16 * nkf * N_PIX  # read once per solve by the pre-pass
    per_solve_bytes = input_bytes + RECORD_BYTES * records + iters * algo_bytes   # + records written once, read per iteration
    roofline = dict(
        bound="hbm", achieved=achieved, peak=peak, unit="GB/s", frac=achieved / peak, traffic=traffic,
        kernel="hessian_kernel<RAYS,HUBER>", peak_source=peak_src, launch_ms=hess_ms,
        unit_of_work="surviving (pixel, directed edge) record, %d B streamed per record and iteration" % RECORD_BYTES,
        surviving_pixel_edges=records, pixel_edges=pixel_edges, survivor_share=records / max(pixel_edges, 1),
        algorithmic_bytes_per_launch=algo_bytes,
        dram_frac=(traffic / (hess_ms * 1e-3) / 1e9 / peak) if traffic else None,
        traffic_over_algorithmic=(traffic / algo_bytes) if traffic and algo_bytes else None,
        per_solve=dict(bytes=per_solve_bytes, gbs=per_solve_bytes / (ms_per_step * 1e-3) / 1e9,
                       frac=per_solve_bytes / (ms_per_step * 1e-3) / 1e9 / peak,
                       what="(inputs read once + records written once + iterations x records) / step time"),
        reference_layout_equivalent_gbs=REF_LAYOUT_BYTES_PER_PIXEL_EDGE * pixel_edges / (hess_ms * 1e-3) / 1e9,
        note="frac = records x 32 B / launch time / measured peak; traffic = DRAM bytes of one launch from the committed "
             "ncu capture of the same workload (profiles/); reference_layout_equivalent_gbs = the 45 B per pixel-edge of "
             "the reference's input layout (SURVEY 8d) over the same time -- not a roofline fraction, because pixels "
             "that fail the pose-independent gates are dropped once per solve by the pre-pass (inside the timed step)")

    # ---- what the reference copies on the device before every solve (global_opt.py:104-119: torch.cat of both edge
    #      directions, torch.stack of every keyframe's pointmap / confidence); the graph-resident FactorGraph mirror
    #      (mast3r_slam_ysh_b200/global_opt.py) hands out views instead
    graph_prep = None
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if rank == 0 and world == 1:
        m = E_local // 2

        def ref_prep():
            a = torch.cat((idx[:m], idx[m:]), 0)
            b = torch.cat((vm[:m], vm[m:]), 0)
            c = torch.cat((Q[:m], Q[m:]), 0)
            d = torch.stack([Xs[k] for k in range(nkf)])
            e = torch.stack([Cs[k] for k in range(nkf)])
            return a, b, c, d, e

        for _ in range(2):
            ref_prep()
        torch.cuda.synchronize()
        k0.record()
        for _ in range(5):
            out = ref_prep()
        k1.record()
        torch.cuda.synchronize()
        copied = sum(t.numel() * t.element_size() for t in out)
        del out
        graph_prep = dict(reference_style_prep_ms_per_solve=k0.elapsed_time(k1) / 5, bytes_copied=int(copied),
                          resident_graph_prep_ms_per_solve=0.0,
                          what="torch.cat of both edge directions + torch.stack of keyframe buffers vs views")
        torch.cuda.empty_cache()

    # ---- the reference's own CUDA kernels on the same GPU and inputs (oracle/_ref; SURVEY 2b's bar)
    ref_cuda = None
    if rank == 0 and world == 1:
        ref_cuda = reference_cuda_extra(torch, be, syn, dev, g, T0, P)

    # ---- e2e: pinned host buffers -> device -> solve -> poses/dx back to the host, every step
    host_in = [x.cpu().pin_memory() for x in (T0, Xs, Cs, ii, jj, idx, vm, Q)]
    h2d = sum(x.numel() * x.element_size() for x in host_in)
    out_T = torch.empty_like(host_in[0]).pin_memory()
    out_dx = torch.empty((nkf - 1, 7), dtype=torch.float32).pin_memory()
    d2h = out_T.numel() * 4 + out_dx.numel() * 4
    if world > 1:
        all_ii_h, all_jj_h = g["all_ii"].cpu().pin_memory(), g["all_jj"].cpu().pin_memory()
        h2d += 2 * all_ii_h.numel() * 8
    del Xs, Cs, idx, vm, Q, g
    torch.cuda.empty_cache()

    def e2e_step():
        d = [x.to(dev, non_blocking=True) for x in host_in]
        if world == 1:
            (dx,) = be.gauss_newton_rays(*d, *P)
        else:
            (dx,), _ = md.gauss_newton_rays_sharded(*d, all_ii_h.to(dev, non_blocking=True),
                                                    all_jj_h.to(dev, non_blocking=True), *P)
        out_T.copy_(d[0], non_blocking=True)
        out_dx.copy_(dx, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(2):
        e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_steps = max(2, min(args.steps, 5))
    for _ in range(e2e_steps):
        e2e_step()
    e1.record()
    barrier()
    e2e_ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_val = float(e2e_ms.item()) / e2e_steps / max(iters, 1)
    del host_in
    torch.cuda.empty_cache()

    # ---- BASELINE.json configs[4]: fixed 256-keyframe loop-closure graph + 256 pair matches, strong-scaled over the ranks
    configs4 = configs4_extra(torch, dist, be, md, syn, dev, rank, world, P)

    extra = {}
    if rank == 0:
        if world == 1:
            extra = matching_extra(be, syn, dev, torch)
            extra["tracker"] = tracker_extra(syn, dev, torch)
            extra["graph_prep"] = graph_prep
            extra["configs3_calib"] = configs3_extra(torch, be, md, _lib, syn, dev)
            extra["hessian_dense_overlap"] = dense_overlap_extra(torch, md, _lib, syn, dev, P, peak)
            extra["reference_cuda"] = ref_cuda
            extra["cpu_tracker"] = cpu_tracker_baseline()
        extra["configs4_strong_scaling"] = configs4
        extra["hessian_kernel_ms"] = hess_ms
        extra["solve_other_ms_per_iter"] = value - hess_ms
        extra["iterations_executed"] = iters
        lanes = (int(info_keep["info"].cpu()[3]) >> 24) & 15
        segs = (int(info_keep["info"].cpu()[3]) >> 20) & 15
        n_sys = 7 * (nkf - 1)
        extra["linear_solver"] = ("single-CTA shared-memory Cholesky (n = %d)" % n_sys if n_sys <= 224 else
                                  "segmented envelope Cholesky, %d interiors on %d CTAs, %d lanes (n = %d)" % (segs, segs + 1, lanes, n_sys) if segs > 0 else
                                  "single-CTA envelope Cholesky, %d lanes (n = %d)" % (lanes, n_sys) if lanes > 0 else
                                  "dense multi-CTA Cholesky (n = %d)" % n_sys)
        cpu_base = cpu_gn_sample(1, device=dev)[0] if world == 1 else None
        line = dict(metric="gn_solve_ms_per_iter", value=value, unit="ms/iter", n_gpus=world, steps=args.steps,
                    warmup=args.warmup, ms_per_step=ms_per_step, higher_is_better=False, scaling="weak", vs_baseline=None,
                    dtype="f32 (fp64 reduction + Cholesky)", data="synthetic", config=workload_config(world),
                    clocks=clocks, gpu_launches=launches,
                    e2e=dict(value=e2e_val, unit="ms/iter", h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=int(d2h)),
                    roofline=roofline, extra=extra)
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def reference_cuda_extra(torch, be, syn, dev, g2, T0, P):
    """The reference's gn_kernels.cu / matching_kernels.cu compiled for sm_100a (oracle/build_ref.py -> oracle/_ref,
    test infrastructure) timed with CUDA events on the same B200 and the same inputs as ours: GN configs[2] and
    configs[3], iter_proj and refine_matches at b = 1."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if ref_dir not in sys.path:
        sys.path.insert(0, ref_dir)
    try:
        import mast3r_slam_backends_ref as ref
    except Exception as ex:  # noqa
        return dict(unavailable=repr(ex))
    lo, mt = syn.LOCAL_OPT, syn.MATCHING
    out = dict(what="reference CUDA (gn_kernels.cu, matching_kernels.cu, Eigen host solve) built for sm_100a, same GPU, "
                    "same inputs, CUDA events; ms per GN iteration = solve time / 10 iterations")
    a2 = [g2[k] for k in ("Xs", "Cs", "ii", "jj", "idx_ii2jj", "valid_match", "Q")]

    def both(name, run_ref, run_ours, reps, per):
        r = _event_time(torch, run_ref, reps, warm=1) / per
        o = _event_time(torch, run_ours, reps, warm=1) / per
        out[name] = dict(reference_ms=r, ours_ms=o, speedup=r / o)

    both("gn_rays_configs2_ms_per_iter", lambda: ref.gauss_newton_rays(T0.clone(), *a2, *P),
         lambda: be.gauss_newton_rays(T0.clone(), *a2, *P), 3, GN_ITERS)
    g3 = syn.make_gn_problem(10, H, W, syn.ring_edges(10, 3, wrap=False), seed=1237, device=dev, calib=True,
                             yaw_step_deg=10.0)
    a3 = [g3[k] for k in ("Xs", "Cs")]
    r3 = [g3[k] for k in ("ii", "jj", "idx_ii2jj", "valid_match", "Q")]
    c3 = (H, W, lo["pixel_border"], lo["depth_eps"], lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"], lo["Q_conf"],
          GN_ITERS, lo["delta_norm"])
    both("gn_calib_configs3_ms_per_iter", lambda: ref.gauss_newton_calib(g3["Twc"].clone(), *a3, g3["K"], *r3, *c3),
         lambda: be.gauss_newton_calib(g3["Twc"].clone(), *a3, g3["K"], *r3, *c3), 5, GN_ITERS)
    del g3, a3, r3
    m = syn.make_match_problem(1, H, W, seed=1235, device=dev)
    rays, pts, p0 = m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"]
    ia = (mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
    both("iter_proj_b1_ms", lambda: ref.iter_proj(rays, pts, p0, *ia), lambda: be.iter_proj(rays, pts, p0, *ia), 20, 1)
    D11, D21 = m["D11"].half(), m["D21"].view(1, N_PIX, -1).half()
    p1 = be.iter_proj(rays, pts, p0, *ia)[0].long()
    both("refine_matches_b1_ms", lambda: ref.refine_matches(D11, D21, p1, mt["radius"], mt["dilation_max"]),
         lambda: be.refine_matches(D11, D21, p1, mt["radius"], mt["dilation_max"]), 20, 1)
    return out


def configs3_extra(torch, be, md, _lib, syn, dev):
    """BASELINE.json configs[3]: FactorGraph.solve_GN_calib, 10 keyframes, edges to the previous 3 -> E = 48, pinhole K,
    points pre-constrained to the pixel rays (gn.cpp:54-82; global_opt.py:160-213)."""
    lo = syn.LOCAL_OPT
    g = syn.make_gn_problem(10, H, W, syn.ring_edges(10, 3, wrap=False), seed=1237, device=dev, calib=True,
                            yaw_step_deg=10.0)
    a = [g[k] for k in ("Xs", "Cs")]
    r = [g[k] for k in ("ii", "jj", "idx_ii2jj", "valid_match", "Q")]
    c = (H, W, lo["pixel_border"], lo["depth_eps"], lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"], lo["Q_conf"],
         GN_ITERS, lo["delta_norm"])
    info = {}

    def run():
        T = g["Twc"].clone()
        (_,), info["i"] = be.gauss_newton_calib(T, *a, g["K"], *r, *c, return_info=True)
        info["T"] = T
    ms = _event_time(torch, run, 10)
    iters = int(info["i"].cpu()[0])
    params = _lib.GnParams(_lib.M3R_GN_CALIB, _lib.M3R_ROBUST_HUBER, lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"],
                           lo["Q_conf"], g["K"].data_ptr(), H, W, lo["pixel_border"], lo["depth_eps"], GN_ITERS,
                           lo["delta_norm"])
    hess_ms, recs = hessian_point(torch, md, _lib, None, g, kind="calib", extra_params=params)
    peak, _ = measured_peak()
    return dict(workload="solve_GN_calib: 10 keyframes, 48 directed edges, 512x384, max_iter 10", solve_ms=ms,
                iterations_executed=iters, ms_per_iter=ms / max(iters, 1), hessian_kernel_ms=hess_ms,
                surviving_pixel_edges=recs, pixel_edges=48 * N_PIX,
                hessian_hbm_frac=RECORD_BYTES * recs / (hess_ms * 1e-3) / 1e9 / peak,
                gt_err_init=float((g["Twc"] - g["Twc_gt"]).abs().max()),
                gt_err_final=float((info["T"] - g["Twc_gt"]).abs().max()))


def dense_overlap_extra(torch, md, _lib, syn, dev, P, peak):
    """The Hessian kernel when the compaction does not thin the work: configs[2]'s graph with a 2 deg yaw step (every
    edge sees nearly the whole image), so that most pixels that pass the Q gate survive."""
    nkf = KF_PER_GPU
    g = syn.make_gn_problem(nkf, H, W, syn.ring_edges(nkf, SPAN), seed=1236, device=dev, yaw_step_deg=2.0)
    ms, recs = hessian_point(torch, md, _lib, P, g, reps=10)
    return dict(workload="configs[2] graph at a 2 deg yaw step", hessian_kernel_ms=ms, surviving_pixel_edges=recs,
                pixel_edges=int(g["ii"].numel()) * N_PIX, survivor_share=recs / (int(g["ii"].numel()) * N_PIX),
                hbm_frac=RECORD_BYTES * recs / (ms * 1e-3) / 1e9 / peak,
                records_per_second=recs / (ms * 1e-3))


def configs4_extra(torch, dist, be, md, syn, dev, rank, world, P):
    """BASELINE.json configs[4] as written: a FIXED 256-keyframe loop-closure graph (consecutive + 3 loop-closure edges
    per keyframe -> 1024 graph edges -> E = 2048 directed, synthetic.loop_closure_edges) solved by gauss_newton_rays
    with the edges sharded over the N ranks, and 256 keyframe-pair matches (iter_proj + refine_matches, and the fused
    match) split by pair.  Strong scaling: total work is the same at every N; times are device times, max over ranks."""
    from mast3r_slam_ysh_b200 import matching as mm
    nkf, n_pairs = 256, 256
    edges = syn.loop_closure_edges(nkf, n_loop=3, seed=1238)
    n_graph = len(edges[0])
    lo_e, hi_e = md.partition_edges(n_graph, world)[rank]
    sel = md.directed_slice(n_graph, lo_e, hi_e)
    g = syn.make_gn_problem(nkf, H, W, edges, seed=1238, device=dev, directed_subset=sel if world > 1 else None,
                            yaw_step_deg=360.0 / nkf)
    a = [g[k] for k in ("Xs", "Cs", "ii", "jj", "idx_ii2jj", "valid_match", "Q")]
    T0 = g["Twc"].clone()
    keep = {}

    def step():
        T = T0.clone()
        if world == 1:
            (_,), keep["info"] = be.gauss_newton_rays(T, *a, *P, return_info=True)
        else:
            (_,), keep["info"] = md.gauss_newton_rays_sharded(T, *a, g["all_ii"], g["all_jj"], *P)
        keep["T"] = T

    def timed(fn, reps, warm):
        for _ in range(warm):
            fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ms = timed(step, 3, 2)
    info = keep["info"].cpu()
    iters = int(info[0])
    out = dict(workload="256-keyframe loop-closure graph, %d graph edges -> %d directed edges, 512x384, max_iter 10; "
                        "256 pair matches" % (n_graph, 2 * n_graph), n_gpus=world,
               gn_solve_ms=ms, gn_iterations_executed=iters, gn_ms_per_iter=ms / max(iters, 1),
               gn_failed_factorisations=int(info[2]), solver_lanes=(int(info[3]) >> 24) & 15,
               gn_gt_err_init=float((T0 - g["Twc_gt"]).abs().max()),
               gn_gt_err_final=float((keep["T"] - g["Twc_gt"]).abs().max()))
    del g, a, keep
    torch.cuda.empty_cache()
    # ---- 256 pair matches split by pair: 8 distinct synthetic pairs repeated to this rank's share
    b_lo, b_hi = md.partition_pairs(n_pairs, world)[rank]
    b_loc = b_hi - b_lo
    m8 = syn.make_match_problem(8, H, W, seed=1239, device=dev)
    rep = lambda t: t.repeat((b_loc + 7) // 8, *([1] * (t.dim() - 1)))[:b_loc].contiguous()
    rays, pts, p0 = rep(m8["rays_img_with_grad"]), rep(m8["pts_3d_norm"]), rep(m8["p_init"])
    D11h, D21h = rep(m8["D11"]).half(), rep(m8["D21"]).view(b_loc, N_PIX, -1).half()
    mt = syn.MATCHING
    t_unf = timed(lambda: md.match_pairs_sharded(rays, pts, p0, D11h, D21h, mt), 3, 1)
    t_ip = timed(lambda: be.iter_proj(rays, pts, p0, mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"]), 3, 1)
    del rays, pts, p0, D11h, D21h
    torch.cuda.empty_cache()
    X11, X21, D11, D21 = rep(m8["X11"]), rep(m8["X21"]), rep(m8["D11"]), rep(m8["D21"])
    t_fus = timed(lambda: mm.match(X11, X21, D11, D21), 3, 1)
    del X11, X21, D11, D21, m8
    torch.cuda.empty_cache()
    out.update(pairs=n_pairs, pairs_per_rank=b_loc,
               iter_proj_ms=t_ip, iter_proj_mpix_s=n_pairs * N_PIX / (t_ip * 1e-3) / 1e6,
               iter_proj_plus_refine_ms=t_unf, iter_proj_plus_refine_mpix_s=n_pairs * N_PIX / (t_unf * 1e-3) / 1e6,
               match_fused_ms=t_fus, match_fused_mpix_s=n_pairs * N_PIX / (t_fus * 1e-3) / 1e6)
    return out


def ncu_traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu capture (profiles/), else null."""
    p = os.path.join(ROOT, "profiles", "hessian_dram_bytes_per_launch.json")
    if os.path.exists(p):
        try:
            return json.load(open(p))["dram_bytes_per_launch"]
        except Exception:
            return None
    return None


def matching_extra(be, syn, dev, torch):
    """BASELINE.json's second metric: iter_proj (+ refine_matches, + the whole match()) Mpix/s at 512x384, configs[1]
    (b = 1) and b = 8, 64, 256 (the bandwidth regime); b > 8 repeats 8 distinct synthetic pairs."""
    from mast3r_slam_ysh_b200 import matching as mm
    mt = syn.MATCHING
    out = {}
    m8 = syn.make_match_problem(8, H, W, seed=1235, device=dev)
    peak, _ = measured_peak()
    for b in (1, 8, 64, 256):
        rep = lambda t: t.repeat((b + 7) // 8, *([1] * (t.dim() - 1)))[:b].contiguous()
        rays, pts, p0 = rep(m8["rays_img_with_grad"]), rep(m8["pts_3d_norm"]), rep(m8["p_init"])
        D11, D21 = rep(m8["D11"]).half(), rep(m8["D21"]).view(b, N_PIX, -1).half()
        fn_ip = lambda: be.iter_proj(rays, pts, p0, mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
        p1 = fn_ip()[0].long()
        fn_rf = lambda: be.refine_matches(D11, D21, p1, mt["radius"], mt["dilation_max"])
        fns = [("iter_proj", fn_ip), ("refine_matches", fn_rf)]
        if b <= 64:
            # whole match() of matching.py:8-91: fused device path (m3r_match) vs the reference's torch glue around the
            # two drop-in kernels
            margs = (rep(m8["X11"]), rep(m8["X21"]), rep(m8["D11"]), rep(m8["D21"]))
            fns += [("match_fused", lambda: mm.match(*margs)), ("match_unfused_torch_glue", lambda: mm.match_unfused(*margs))]
        res = {}
        for name, fn in fns:
            reps = 20 if b <= 8 else 5
            ms = _event_time(torch, fn, reps, warm=2)
            res[name + "_ms"] = ms
            res[name + "_mpix_s"] = b * N_PIX / (ms * 1e-3) / 1e6
        res["iter_proj_hbm_frac"] = ITER_PROJ_BYTES_PER_QUERY * b * N_PIX / (res["iter_proj_ms"] * 1e-3) / 1e9 / peak
        out[f"matching_b{b}"] = res
        del rays, pts, p0, D11, D21, p1, fns
        margs = None
        torch.cuda.empty_cache()
    return out


def tracker_extra(syn, dev, torch):
    """BASELINE.json configs[0] workload (one 512x384 pair, FrameTracker.opt_pose_ray_dist_sim3 with the thresholds of
    config/base.yaml) on the device tracker kernel (m3r_track_rays), next to the reference's own
    formulation run here: the torch-eager restatement (oracle/tracker_torch.py) on the same GPU -- what the reference's
    tracker executes on a CUDA device -- and on the host cores (the CPU baseline of configs[0])."""
    from mast3r_slam_ysh_b200 import tracker as trk
    from oracle import tracker_torch as tt
    t = syn.make_track_problem(H, W, seed=1234, device="cpu")
    d = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in t.items()}
    fn = lambda: trk.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Qk"], d["valid"], check=False,
                                            return_info=True)
    for _ in range(3):
        info = fn()[2]
    torch.cuda.synchronize()
    a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    a.record()
    for _ in range(reps):
        fn()
    c.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(c) / reps
    its = info.iterations
    # the reference's formulation, eager torch on the same GPU (a .item() sync per iteration, as tracker.py:166)
    eager = []
    for rep in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _, _, its_e = tt.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], tt.Sim3(d["T_WCf"].clone()), tt.Sim3(d["T_WCk"].clone()),
                                                d["Qk"], d["valid"])
        torch.cuda.synchronize()
        if rep:
            eager.append((time.perf_counter() - t0) * 1e3)
    eager.sort()
    return dict(device_solve_ms=ms, device_iters=its, device_ms_per_iter=ms / max(its, 1),
                torch_eager_gpu_solve_ms=eager[len(eager) // 2], torch_eager_gpu_iters=its_e,
                what="tracker.opt_pose_ray_dist_sim3, 1 pair 512x384, config/base.yaml thresholds")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
