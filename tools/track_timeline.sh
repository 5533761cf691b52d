#!/bin/bash
# dev tool (GPU box): per-stage clock counts of the tracker kernel's first iterations, then the production timing
cd mast3r_slam_ysh_b200/csrc
rm -f _obj/track.o; make -s EXTRA="-DM3R_TRK_TIMELINE" > /dev/null 2>&1 || echo "variant build failed"
(cd ../..; python - <<'PY'
import torch
from mast3r_slam_ysh_b200 import synthetic as syn, tracker as trk
t = syn.make_track_problem(384, 512, seed=1234)
d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in t.items()}
cfg = dict(max_iters=6, rel_error=0.0, delta_norm=0.0)
for _ in range(2):
    trk.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Qk"], d["valid"], cfg=cfg)
    torch.cuda.synchronize()
    print("--")
PY
)
rm -f _obj/track.o; make -s > /dev/null 2>&1
cd ../..; python tools/time_track.py
