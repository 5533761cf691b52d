"""GPU tests of the drop-in boundary: the compiled pybind module and the ctypes mirror are two host bindings of the same
C-ABI library and must agree bit for bit; the host-buffer entry points of the C ABI (m3r_*_host: plain host pointers
in, results back on the host) reproduce the device entry points; per-call host overhead of both bindings is recorded."""
import ctypes as C
import importlib
import json
import os
import time

import numpy as np
import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _syn():
    from mast3r_slam_ysh_b200 import synthetic as syn
    return syn


def test_compiled_binding_equals_ctypes_mirror(be):
    syn = _syn()
    ext = importlib.import_module("mast3r_slam_ysh_b200.mast3r_slam_backends_b200")
    lo, mt = syn.LOCAL_OPT, syn.MATCHING
    m = syn.make_match_problem(2, 96, 128, seed=9, device=DEV)
    ia = (m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"], mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
    p_a, c_a = be.iter_proj(*ia)
    p_b, c_b = ext.iter_proj(*ia)
    assert torch.equal(p_a, p_b) and torch.equal(c_a, c_b) and c_b.dtype == torch.bool
    D11, D21 = m["D11"].half(), m["D21"].view(2, 96 * 128, -1).half()
    (q_a,) = be.refine_matches(D11, D21, p_a.long(), mt["radius"], mt["dilation_max"])
    out = ext.refine_matches(D11, D21, p_a.long(), mt["radius"], mt["dilation_max"])
    assert isinstance(out, list) and len(out) == 1 and torch.equal(out[0], q_a)
    for kind in ("rays", "calib", "points"):
        g = syn.make_gn_problem(6, 48, 64, syn.ring_edges(6, 2, wrap=False), seed=5, device=DEV, calib=(kind == "calib"),
                                yaw_step_deg=9.0, kf_ids=[2, 3, 5, 8, 9, 11])
        a = [g[k] for k in ("Xs", "Cs")]
        r = [g[k] for k in ("ii", "jj", "idx_ii2jj", "valid_match", "Q")]
        Ta, Tb = g["Twc"].clone(), g["Twc"].clone()
        if kind == "rays":
            tail = (lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], 4, lo["delta_norm"])
            (da,) = be.gauss_newton_rays(Ta, *a, *r, *tail)
            (db,) = ext.gauss_newton_rays(Tb, *a, *r, *tail)
        elif kind == "points":
            tail = (lo["sigma_point"], lo["C_conf"], lo["Q_conf"], 4, lo["delta_norm"])
            (da,) = be.gauss_newton_points(Ta, *a, *r, *tail)
            (db,) = ext.gauss_newton_points(Tb, *a, *r, *tail)
        else:
            tail = (48, 64, lo["pixel_border"], lo["depth_eps"], lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"],
                    lo["Q_conf"], 4, lo["delta_norm"])
            (da,) = be.gauss_newton_calib(Ta, *a, g["K"], *r, *tail)
            (db,) = ext.gauss_newton_calib(Tb, *a, g["K"], *r, *tail)
        torch.cuda.synchronize()
        assert torch.equal(Ta, Tb) and torch.equal(da, db), kind   # Twc updated in place by both
        assert not torch.equal(Ta, g["Twc"])
    with pytest.raises(RuntimeError, match="contiguous"):  # CHECK_CONTIGUOUS, gn.h:5
        ext.iter_proj(m["rays_img_with_grad"].transpose(1, 2), m["pts_3d_norm"], m["p_init"], 10, 1e-8, 1e-6)


def test_host_buffer_entry_points_of_the_c_abi(be):
    """m3r_iter_proj_host / m3r_refine_matches_f16_host / m3r_gauss_newton_host: host pointers in and out."""
    from mast3r_slam_ysh_b200 import _lib
    syn = _syn()
    lib = _lib.load()
    lo, mt = syn.LOCAL_OPT, syn.MATCHING
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    b, h, w = 2, 40, 56
    m = syn.make_match_problem(b, h, w, seed=4)
    rays, pts, p0 = (np.ascontiguousarray(m[k].numpy()) for k in ("rays_img_with_grad", "pts_3d_norm", "p_init"))
    p_new = np.zeros((b, h * w, 2), np.float32)
    conv = np.zeros((b, h * w), np.uint8)
    rc = lib.m3r_iter_proj_host(P(rays), P(pts), P(p0), P(p_new), P(conv), b, h, w, h * w, mt["max_iter"], mt["lambda_init"],
                                mt["convergence_thresh"])
    assert rc == 0
    p_d, c_d = be.iter_proj(m["rays_img_with_grad"].to(DEV), m["pts_3d_norm"].to(DEV), m["p_init"].to(DEV), mt["max_iter"],
                            mt["lambda_init"], mt["convergence_thresh"])
    assert np.array_equal(p_new, p_d.cpu().numpy()) and np.array_equal(conv.astype(bool), c_d.cpu().numpy())
    D11 = np.ascontiguousarray(m["D11"].half().numpy()).view(np.uint16)
    D21 = np.ascontiguousarray(m["D21"].view(b, h * w, -1).half().numpy()).view(np.uint16)
    p1 = np.ascontiguousarray(p_d.long().cpu().numpy())
    q = np.zeros_like(p1)
    rc = lib.m3r_refine_matches_f16_host(P(D11), P(D21), P(p1), P(q), b, h, w, h * w, 24, mt["radius"], mt["dilation_max"])
    assert rc == 0
    (q_d,) = be.refine_matches(m["D11"].half().to(DEV), m["D21"].view(b, h * w, -1).half().to(DEV), p_d.long(), mt["radius"],
                               mt["dilation_max"])
    assert np.array_equal(q, q_d.cpu().numpy())
    for kind in ("rays", "calib"):
        g = syn.make_gn_problem(5, 32, 48, syn.ring_edges(5, 2, wrap=False), seed=8, calib=(kind == "calib"), yaw_step_deg=9.0)
        n = {k: np.ascontiguousarray(v.numpy()) for k, v in g.items() if isinstance(v, torch.Tensor)}
        T = n["Twc"].copy()
        dx = np.zeros((4, 7), np.float32)
        info = np.zeros(4, np.int32)
        if kind == "rays":
            params = _lib.GnParams(_lib.M3R_GN_RAYS, 0, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], None, 0,
                                   0, 0, 0.0, 5, lo["delta_norm"])
        else:
            params = _lib.GnParams(_lib.M3R_GN_CALIB, 0, lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"], lo["Q_conf"],
                                   n["K"].ctypes.data, 32, 48, lo["pixel_border"], lo["depth_eps"], 5, lo["delta_norm"])
        rc = lib.m3r_gauss_newton_host(C.byref(params), P(T), P(n["Xs"]), P(n["Cs"]), P(n["ii"]), P(n["jj"]), P(n["idx_ii2jj"]),
                                       P(n["valid_match"].view(np.uint8)), P(n["Q"]), 5, 32 * 48, int(n["ii"].shape[0]), P(dx),
                                       P(info))
        assert rc == 0, lib.m3r_last_error_string()
        Td = g["Twc"].to(DEV)
        d = [g[k].to(DEV) for k in ("Xs", "Cs", "ii", "jj", "idx_ii2jj", "valid_match", "Q")]
        if kind == "rays":
            (dxd,), infod = be.gauss_newton_rays(Td, *d, lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], 5,
                                                 lo["delta_norm"], return_info=True)
        else:
            (dxd,), infod = be.gauss_newton_calib(Td, d[0], d[1], g["K"].to(DEV), *d[2:], 32, 48, lo["pixel_border"],
                                                  lo["depth_eps"], lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"],
                                                  lo["Q_conf"], 5, lo["delta_norm"], return_info=True)
        assert np.array_equal(T, Td.cpu().numpy()) and np.array_equal(dx, dxd.cpu().numpy()), kind
        assert info[0] == 5 and np.array_equal(info, infod.cpu().numpy())
        assert not np.array_equal(T, n["Twc"])


def test_per_call_host_overhead_of_both_bindings(be):
    """Host time per call on problems small enough that the device work is negligible (what the tracker-rate callers of
    the reference pay per frame): recorded under gpurun_out/binding_overhead.json, loose sanity bound only."""
    syn = _syn()
    ext = importlib.import_module("mast3r_slam_ysh_b200.mast3r_slam_backends_b200")
    mt, lo = syn.MATCHING, syn.LOCAL_OPT
    m = syn.make_match_problem(1, 16, 24, seed=2, device=DEV)
    ia = (m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"], mt["max_iter"], mt["lambda_init"], mt["convergence_thresh"])
    g = syn.make_gn_problem(3, 16, 24, ([0, 1], [1, 2]), seed=3, device=DEV, yaw_step_deg=6.0)
    ga = [g[k] for k in ("Xs", "Cs", "ii", "jj", "idx_ii2jj", "valid_match", "Q")]
    gt = (lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], 1, lo["delta_norm"])
    res = {}
    for name, mod in (("ctypes_mirror", be), ("compiled_pybind", ext)):
        for fn_name, fn in (("iter_proj", lambda: mod.iter_proj(*ia)),
                            ("gauss_newton_rays_1iter", lambda: mod.gauss_newton_rays(g["Twc"].clone(), *ga, *gt))):
            for _ in range(20):
                fn()
            torch.cuda.synchronize()
            t = time.perf_counter()
            reps = 200
            for _ in range(reps):
                fn()
            host = (time.perf_counter() - t) / reps   # enqueue only: no synchronisation inside the loop
            torch.cuda.synchronize()
            res[f"{name}.{fn_name}_host_us_per_call"] = host * 1e6
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "binding_overhead.json"), "w"), indent=1)
    assert all(v < 5000 for v in res.values()), res
