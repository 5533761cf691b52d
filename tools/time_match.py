"""Fused m3r_match vs the reference's torch glue around the drop-in kernels (match_unfused), 512x384.  CUDA events."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import matching as mm, synthetic as syn
dev = "cuda:0"; out = {}
for b in (1, 8):
    m = syn.make_match_problem(b, 384, 512, seed=1235, device=dev)
    a = (m["X11"], m["X21"], m["D11"], m["D21"])
    def t(fn, reps=10):
        for _ in range(3): fn()
        torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); e0.record()
        for _ in range(reps): fn()
        e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / reps
    f = t(lambda: mm.match(*a)); u = t(lambda: mm.match_unfused(*a))
    i_f, v_f = mm.match(*a); i_u, v_u = mm.match_unfused(*a)
    out[f"b{b}"] = dict(fused_ms=f, unfused_ms=u, same_idx=float((i_f == i_u).float().mean()), same_valid=float((v_f == v_u).float().mean()),
                        valid_frac=float(v_f.float().mean()), mpix_s_fused=b * 384 * 512 / f / 1e3)
print(json.dumps(out))
