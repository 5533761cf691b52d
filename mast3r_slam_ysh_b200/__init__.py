"""B200-native (sm_100a) dense backend of MASt3R-SLAM: drop-in for the reference's `mast3r_slam_backends`.

Layout: csrc/ (CUDA kernels + C ABI -> libm3r_backends_b200.so), _lib.py (ctypes binding), backends.py (the
reference's five-function operator interface on torch tensors), distributed.py (edge-/pair-sharded multi-GPU
drivers), synthetic.py (input generators for BASELINE.json's configs).
"""
from . import backends  # noqa: F401
from .backends import (gauss_newton_calib, gauss_newton_points, gauss_newton_rays, iter_proj,  # noqa: F401
                       refine_matches)

__version__ = "0.1.0"
