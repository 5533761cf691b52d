"""Drop-in module name of the reference extension (`import mast3r_slam_backends`, matching.py:5, global_opt.py:9).

Re-exports the five entry points of gn.cpp:116-123 from the B200-native implementation.
"""
from mast3r_slam_ysh_b200.backends import (gauss_newton_calib, gauss_newton_points, gauss_newton_rays,  # noqa: F401
                                           iter_proj, refine_matches)
