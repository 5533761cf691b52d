"""Phase timeline of the fused solver kernel (m3r_debug_set_timeline): prints the time CTA 0 spends in each phase of one
warm GN iteration.  Usage: python tools/fused_timeline.py [nkf] [span]   (span 0 = loop-closure graph)"""
import os, sys, json, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from mast3r_slam_ysh_b200 import _lib, distributed as md, synthetic as syn
nkf = int(sys.argv[1]) if len(sys.argv) > 1 else 32
span = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda:0"); lo = syn.LOCAL_OPT
edges = syn.ring_edges(nkf, span) if span > 0 else syn.loop_closure_edges(nkf)
g = syn.make_gn_problem(nkf, 384, 512, edges, seed=1236, device=dev)
P = _lib.GnParams(_lib.M3R_GN_RAYS, 0, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], None, 0, 0, 0, 0.0, 10, 1e-8)
T = g["Twc"].clone()
_lib.load().m3r_debug_set_solver(int(os.environ.get("M3R_SOLVER_MODE", "0")))  # the plan is made by gn_begin
s = md.ShardedGaussNewton(P, T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], g["all_ii"], g["all_jj"])
lib = s.lib
for _ in range(3): s.iterate()
torch.cuda.synchronize()
cap = 4096
buf = torch.zeros(2 * cap, dtype=torch.int64, device=dev)
lib.m3r_debug_set_timeline(C.c_void_p(buf.data_ptr()), cap)
if os.environ.get("M3R_TL_SYNTH"):
    # experiment: same launch, benign matrix (is the factorisation time data dependent?)
    rc = lib.m3r_gn_accumulate(C.byref(P), C.c_void_p(T.data_ptr()), s.num_kf, s.num_pts, s.E, C.c_void_p(s.ws_ptr), s._stream())
    torch.cuda.synchronize()
    sysv = s.sys
    print("sys stats: absmax %.3e, min nonzero abs %.3e, nonfinite %d" % (float(sysv.abs().max()), float(sysv[sysv != 0].abs().min()), int((~torch.isfinite(sysv)).sum())))
    if os.environ["M3R_TL_SYNTH"] == "2":
        sysv.mul_(1e-6)
    rc = lib.m3r_gn_solve_retract(C.byref(P), C.c_void_p(T.data_ptr()), s.num_kf, s.num_pts, s.E, C.c_void_p(s.ws_ptr), s._stream())
else:
    s.iterate()
torch.cuda.synchronize()
lib.m3r_debug_set_timeline(None, 0)
b = buf.cpu().tolist(); k = b[0]
marks = [(b[2 * (i + 1)], b[2 * (i + 1) + 1]) for i in range(k)]
names = {1: "start", 2: "finalize+zero", 3: "barrier", 4: "assemble", 5: "barrier/start-factor", 6: "backsolve|finish", 7: "small_backsolve", 8: "small_load",
         8100: "nd_interiors_eliminated(wait)", 8200: "nd_reduce", 8300: "nd_separator_factor", 8400: "nd_separator_backsolve", 8500: "nd_interiors_backsolve(wait)"}
agg = {}
small = 7 * (nkf - 1) <= 224
prev = None
cyc = [(tag - 10000, t) for tag, t in marks if tag >= 10000]
marks = [m for m in marks if m[0] < 9000]
sky_marks = [m for m in marks if 7000 <= m[0] < 8000]  # steps 20/21 of the envelope solver (SKY_MARK)
marks = [m for m in marks if not (7000 <= m[0] < 8000)]
if not small: names[8] = "envelope_factor"
if any(5000 <= tag < 6000 for tag, _ in marks):  # -DM3R_PANEL_MARKS: steps inside panel 3 of CTA 0
    pm = [(tag, t) for tag, t in marks if tag < 5000 or tag in (5100, 5200, 5300, 5400, 5500)]
    for i, (tag, t) in enumerate(pm):
        if tag == 5100:
            seq = pm[i - 1:i + 6]
            print("panel 3 steps (us): " + ", ".join("%d:%.2f" % (seq[j + 1][0], (seq[j + 1][1] - seq[j][1]) / 1e3) for j in range(len(seq) - 1)))
    marks = [m for m in marks if not (5000 <= m[0] < 6000)]
for tag, t in marks:
    if tag == 1: prev = t; continue
    key = names.get(tag) or ({1: "first_factor(kb=0 only)", 2: "trsm", 3: "diagupd+factor_next||update"} if small else {1: "panel", 2: "barrier_after_panel", 3: "update", 4: "barrier_after_update"})[tag // 1000]
    agg.setdefault(key, []).append((t - prev) / 1e3); prev = t
out = {"nkf": nkf, "n": 7 * (nkf - 1), "marks": k, "solver_mode": int(os.environ.get("M3R_SOLVER_MODE", "0")),
       "segments": int(lib.m3r_debug_solver_segments(C.c_void_p(s.ws_ptr)))}
for key, v in agg.items():
    out[key] = {"count": len(v), "sum_us": round(sum(v), 2), "max_us": round(max(v), 2), "first_us": round(v[0], 2)}
if cyc:  # clock64 stamps inside the single-CTA solver, relative to its start (1.965 GHz)
    nm = {1: "start", 8: "loaded", 7: "backsolved"}
    prevc, agg2 = 0, {}
    for tag, c in cyc:
        key = nm.get(tag) or {1: "factor0|loopstart", 2: "trsm", 3: "diagupd+factor_next||update"}[tag // 1000]
        agg2.setdefault(key, []).append(c - prevc); prevc = c
    out["small_solver_cycles"] = {k: {"sum": sum(v), "max": max(v), "n": len(v)} for k, v in agg2.items()}
    out["small_solver_total_us"] = round(prevc / 1965.0, 2)
if sky_marks:
    t0 = sky_marks[0][1]
    out["envelope_step_20_21_us"] = [(tag, round((t - t0) / 1e3, 2)) for tag, t in sky_marks]
print(json.dumps(out))
