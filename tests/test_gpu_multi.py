"""Multi-GPU tests (-m gpu; skipped unless the box has at least two devices -- run with `gpurun --gpus 2`).

One torchrun job over min(4, device_count) ranks drives tests/multi_gpu_worker.py: edge-sharded Gauss-Newton (rays, calib,
points; in-kernel peer all-reduce and NCCL) against the single-GPU solve, and pair-sharded matching against the single-GPU
result.  A second test runs the single-process entry points on cuda:1 after cuda:0 (the shared-memory opt-ins of the
kernels are per device)."""
import json
import os
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _ndev():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.skipif(_ndev() < 2, reason="needs at least two CUDA devices")
def test_sharded_gn_and_pair_sharded_matching_vs_single_gpu():
    n = 4 if _ndev() >= 4 else 2
    out = os.path.join(ROOT, "gpurun_out", f"multi_gpu_n{n}.json")
    port = 29600 + os.getpid() % 1000
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "multi_gpu_worker.py"), out]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.load(open(out))
    assert res["ok"] and res["world"] == n
    assert res["match_pairs_sharded"]["unfused_equal"] and res["match_pairs_sharded"]["fused_equal"]


@pytest.mark.skipif(_ndev() < 2, reason="needs at least two CUDA devices")
def test_entry_points_on_second_device_after_first(be):
    """cudaFuncSetAttribute applies to the current device only: a process that used cuda:0 must still be able to launch the
    kernels with large dynamic shared memory (small solver 192 KB, envelope / dense solver, tracker, pack) on cuda:1."""
    from mast3r_slam_ysh_b200 import synthetic as syn, tracker as trk
    lo = syn.LOCAL_OPT
    res = []
    for dev in ("cuda:0", "cuda:1"):
        for nkf, span in ((6, 2), (40, 3)):  # small solver; envelope solver
            g = syn.make_gn_problem(nkf, 24, 32, syn.ring_edges(nkf, span, wrap=False), seed=5, device=dev, yaw_step_deg=7.0)
            T = g["Twc"].clone()
            with torch.cuda.device(dev):
                (dx,), info = be.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"],
                                                   g["Q"], lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], 3,
                                                   lo["delta_norm"], return_info=True)
                torch.cuda.synchronize(dev)
            assert info.cpu()[2] == 0 and dx.device == torch.device(dev)
            res.append(T.cpu())
        t = syn.make_track_problem(48, 64, seed=3, device=dev)
        Tf, Tr, info = trk.opt_pose_ray_dist_sim3(t["Xf"], t["Xk"], t["T_WCf"], t["T_WCk"], t["Qk"], t["valid"],
                                                  return_info=True)
        torch.cuda.synchronize(dev)
        res.append(Tf.cpu())
    k = len(res) // 2
    for a, b in zip(res[:k], res[k:]):
        assert torch.equal(a, b)  # same inputs, same bits on either device
