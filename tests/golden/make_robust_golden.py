"""Generates tests/golden/robust.npz from the REFERENCE's own Python (mast3r_slam/nonlinear_optimizer.py:5-42), imported
in this container from /root/reference (CPU torch; no GPU needed).  Run:  python tests/golden/make_robust_golden.py

Pins huber / tukey / check_convergence for (a) the C oracle's robust weights, (b) the tracker restatement
(oracle/tracker_torch.py) and, through (a), (c) the ROBUST template parameter of the sm_100a Hessian kernel.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

REF = os.environ.get("M3R_REFERENCE_ROOT", "/root/reference")
spec = importlib.util.spec_from_file_location("ref_nlo", os.path.join(REF, "mast3r_slam", "nonlinear_optimizer.py"))
nlo = importlib.util.module_from_spec(spec)
spec.loader.exec_module(nlo)

g = torch.Generator().manual_seed(42)
r = torch.cat((4.0 * torch.randn(4000, generator=g), torch.linspace(-6, 6, 2001),
               torch.tensor([0.0, -0.0, 1.345, -1.345, 1.3449999, 1.3450001, 4.6851, -4.6851, 4.68509, 4.68511, 1e-30, 1e30,
                             float("inf")]))).float()
huber = nlo.huber(r)
tukey = nlo.tukey(r)
# check_convergence(iter, rel_error_threshold, delta_norm_threshold, old_cost, new_cost, delta)
cases, conv = [], []
for old, new, dn in [(10.0, 9.0, 1.0), (10.0, 9.995, 1.0), (10.0, 10.004, 1.0), (10.0, 5.0, 5e-4), (1e-3, 2e-3, 1.0),
                     (7.0, 7.0, 0.5), (3.0, 2.9971, 2e-3), (3.0, 2.9969, 1e-3), (3.0, 3.1, 9.99e-4)]:
    delta = torch.zeros(7)
    delta[2] = dn
    cases.append((old, new, dn))
    conv.append(bool(nlo.check_convergence(0, 1e-3, 1e-3, old, new, delta)))
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "robust.npz")
np.savez_compressed(out, r=r.numpy(), huber=huber.numpy(), tukey=tukey.numpy(), cc_cases=np.array(cases, np.float64),
                    cc_converged=np.array(conv, bool), rel_error=1e-3, delta_norm=1e-3)
print("wrote", out, "huber", float(huber.mean()), "tukey", float(tukey.mean()), "conv", conv)
