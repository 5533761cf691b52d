"""Time the device tracker solve (m3r_track_rays / m3r_track_calib) on one 512x384 pair: ms per call and per iteration
for a fixed iteration count and with the reference's convergence thresholds.  usage: python tools/time_track.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mast3r_slam_ysh_b200 import synthetic as syn  # noqa: E402
from mast3r_slam_ysh_b200 import tracker as trk  # noqa: E402


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    torch.cuda.synchronize()
    ev[0].record()
    for i in range(reps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))
    return ts[len(ts) // 2]


def main():
    out = {}
    for calib in (False, True):
        t = syn.make_track_problem(384, 512, seed=1234, calib=calib)
        d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in t.items()}

        def run(cfg):
            if calib:
                return trk.opt_pose_calib_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Qk"], d["valid"], d["meas_k"],
                                               d["valid_meas_k"], d["K"], d["img_size"], cfg=cfg, check=False,
                                               return_info=True)
            return trk.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Qk"], d["valid"], cfg=cfg,
                                              check=False, return_info=True)

        r = {}
        for iters in (1, 10, 50):
            cfg = dict(max_iters=iters, rel_error=0.0, delta_norm=0.0)
            r[f"fixed_{iters}_ms"] = timed(lambda: run(cfg))
        r["per_iter_us"] = (r["fixed_50_ms"] - r["fixed_10_ms"]) / 40 * 1e3
        r["default_ms"] = timed(lambda: run(None))
        r["default_iters"] = run(None)[2].iterations
        r["valid_frac"] = t["valid"].float().mean().item()
        # GPU time alone: 20 calls captured into one CUDA graph (no host work between the launches)
        for name, cfg in (("graph_default_us", None), ("graph_fixed_10_us", dict(max_iters=10, rel_error=0.0, delta_norm=0.0))):
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            with torch.cuda.stream(side):
                run(cfg)
                torch.cuda.synchronize()
                with torch.cuda.graph(g, stream=side):
                    for _ in range(20):
                        keep = run(cfg)
            r[name] = timed(g.replay, reps=10) / 20 * 1e3
        out["calib" if calib else "rays"] = r
    print(json.dumps(out))


if __name__ == "__main__":
    main()
