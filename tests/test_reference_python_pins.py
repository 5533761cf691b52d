"""CPU tests: the robust kernels and the tracker stop rule are pinned to the REFERENCE's own Python
(mast3r_slam/nonlinear_optimizer.py:5-42) -- through tests/golden/robust.npz (made by tests/golden/make_robust_golden.py
importing that file in this container) and, where the file is reachable (/root/reference here, oracle/_ref/pyref on the
GPU box), by importing it live.

Chain of evidence for the opt-in Tukey kernel: reference Python -> golden -> C oracle
(orc_robust_weights, the function orc_pixel_rows uses) -> sm_100a Hessian kernel (tests/test_gpu_parity.py::
test_edge_blocks_vs_oracle_fp64 runs robust="tukey" against the oracle)."""
import importlib.util
import math
import os

import numpy as np
import pytest
import torch

from conftest import ROOT, load_golden


def _ref_python_root():
    for root in (os.environ.get("M3R_REFERENCE_ROOT", "/root/reference"), os.path.join(ROOT, "oracle", "_ref", "pyref")):
        if os.path.exists(os.path.join(root, "mast3r_slam", "nonlinear_optimizer.py")):
            return root
    return None


def _load_ref_nlo():
    root = _ref_python_root()
    if root is None:
        pytest.skip("reference Python not reachable (neither /root/reference nor oracle/_ref/pyref)")
    spec = importlib.util.spec_from_file_location("ref_nlo", os.path.join(root, "mast3r_slam", "nonlinear_optimizer.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _close(a, b, rtol):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    fin = np.isfinite(a) & np.isfinite(b)
    assert np.array_equal(np.isfinite(a), np.isfinite(b))
    assert np.all(np.abs(a[fin] - b[fin]) <= rtol * np.maximum(np.abs(b[fin]), 1e-30))


def test_oracle_c_huber_and_tukey_match_reference_python_golden(orc):
    g = load_golden("robust")
    r = g["r"]
    # Huber: the CUDA reference compares/divides against the double literal 1.345 (gn_kernels.cu:172-175), the Python
    # one in float32 -> identical up to one rounding of k / |r|
    _close(orc.robust_weights(r, orc.ROBUST_HUBER), g["huber"], 2e-7)
    # Tukey (nonlinear_optimizer.py:36-42): (1 - (|r|/t)^2)^2 inside, 0 outside; one ulp of (r/t)^2 near |r| = t
    got = orc.robust_weights(r, orc.ROBUST_TUKEY)
    want = g["tukey"]
    assert np.array_equal(got == 0.0, want == 0.0) or np.abs(np.abs(r[(got == 0) != (want == 0)]) - 4.6851).max() < 1e-5
    assert np.abs(got - want).max() <= 1e-6
    assert want[np.abs(r) >= 4.69].max() == 0.0 and want[np.abs(r) < 1e-3].min() > 0.999


def test_tracker_restatement_huber_and_stop_rule_match_reference_python_golden():
    from oracle import tracker_torch as tt
    g = load_golden("robust")
    r = torch.from_numpy(g["r"])
    assert np.array_equal(tt.huber(r).numpy(), g["huber"])
    for (old, new, dn), want in zip(g["cc_cases"], g["cc_converged"]):
        delta = torch.zeros(7)
        delta[2] = float(dn)
        got = tt.check_convergence(float(g["rel_error"]), float(g["delta_norm"]), float(old), float(new), delta)
        assert bool(got) == bool(want), (old, new, dn)


def test_golden_equals_live_reference_python():
    """Keeps the fixture honest wherever the reference file can be imported."""
    nlo = _load_ref_nlo()
    g = load_golden("robust")
    r = torch.from_numpy(g["r"])
    assert np.array_equal(nlo.huber(r).numpy(), g["huber"])
    assert np.array_equal(nlo.tukey(r).numpy(), g["tukey"], equal_nan=True)
    for (old, new, dn), want in zip(g["cc_cases"], g["cc_converged"]):
        delta = torch.zeros(7)
        delta[2] = float(dn)
        assert bool(nlo.check_convergence(0, 1e-3, 1e-3, float(old), float(new), delta)) == bool(want)


def test_tracker_restatement_uses_the_reference_stop_rule_first_iteration():
    """tracker.py:166 passes old_cost = inf on the first iteration: |(inf - c) / inf| is NaN, NaN < thr is False."""
    from oracle import tracker_torch as tt
    assert not tt.check_convergence(1e-3, 1e-3, float("inf"), 5.0, torch.ones(7))
    assert tt.check_convergence(1e-3, 1e-3, float("inf"), 5.0, 1e-4 * torch.ones(7))
    assert math.isnan(math.fabs((float("inf") - 5.0) / float("inf")))
