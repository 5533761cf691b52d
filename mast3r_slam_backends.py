"""Drop-in module name of the reference extension (`import mast3r_slam_backends`, matching.py:5, global_opt.py:9).

Re-exports the five entry points of gn.cpp:116-123 from the B200-native implementation.  Two host bindings of the same
C-ABI library (libm3r_backends_b200.so) exist and behave identically:
  * the compiled torch C++ extension mast3r_slam_ysh_b200/mast3r_slam_backends_b200*.so (csrc/ext/gn_b200.cpp: pybind,
    torch::Tensor signatures exactly as gn.h declares them) -- used when it has been built;
  * the ctypes mirror mast3r_slam_ysh_b200/backends.py -- used otherwise, or with M3R_BINDING=ctypes.
Neither has a CPU path.
"""
import os as _os

BINDING = "ctypes"
if _os.environ.get("M3R_BINDING", "auto") != "ctypes":
    try:
        import torch as _torch  # noqa: F401  (libtorch must be loaded before the extension)
        from mast3r_slam_ysh_b200.mast3r_slam_backends_b200 import (gauss_newton_calib, gauss_newton_points,  # noqa: F401
                                                                   gauss_newton_rays, iter_proj, refine_matches)
        BINDING = "compiled"
    except ImportError:
        if _os.environ.get("M3R_BINDING") == "compiled":
            raise
if BINDING == "ctypes":
    from mast3r_slam_ysh_b200.backends import (gauss_newton_calib, gauss_newton_points, gauss_newton_rays,  # noqa: F401
                                               iter_proj, refine_matches)
