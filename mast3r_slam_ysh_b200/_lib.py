"""ctypes binding of libm3r_backends_b200.so (the C ABI declared in include/m3r_backends.h).

There is NO fallback: if the CUDA library is missing or fails to load, importing the product path raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# M3R_LIB_PATH: development switch for timing compile-time variants of the library (tools/build_variants.sh)
LIB_PATH = os.environ.get("M3R_LIB_PATH") or os.path.join(_HERE, "libm3r_backends_b200.so")

M3R_GN_POINTS, M3R_GN_RAYS, M3R_GN_CALIB = 0, 1, 2
M3R_ROBUST_HUBER, M3R_ROBUST_TUKEY = 0, 1


class GnParams(C.Structure):
    """struct m3r_gn_params (include/m3r_backends.h)"""
    _fields_ = [
        ("kind", C.c_int32),
        ("robust", C.c_int32),
        ("sigma_a", C.c_float),
        ("sigma_b", C.c_float),
        ("C_thresh", C.c_float),
        ("Q_thresh", C.c_float),
        ("K", C.c_void_p),
        ("height", C.c_int32),
        ("width", C.c_int32),
        ("pixel_border", C.c_int32),
        ("z_eps", C.c_float),
        ("max_iter", C.c_int32),
        ("delta_thresh", C.c_float),
    ]


# name -> (restype, argtypes); every symbol include/m3r_backends.h declares
_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
_ll, _d = C.c_longlong, C.c_double
_PP = C.POINTER(GnParams)
SYMBOLS = {
    "m3r_abi_version": (_i, []),
    "m3r_last_error_string": (C.c_char_p, []),
    "m3r_launch_count": (C.c_uint64, []),
    "m3r_debug_set_timeline": (None, [_vp, _i]),
    "m3r_debug_set_solver": (None, [_i]),
    "m3r_debug_set_refine_mode": (None, [_i]),
    "m3r_debug_set_iter_proj_mode": (None, [_i]),
    "m3r_debug_set_iter_proj_stats": (None, [_vp]),
    "m3r_debug_solver_lanes": (_i, [_vp]),
    "m3r_debug_solver_segments": (_i, [_vp]),
    "m3r_debug_nd_plan_ints": (_i, []),
    "m3r_debug_nd_plan": (_i, [_vp, _i, _i, _i, _vp]),
    "m3r_iter_proj": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _f, _vp]),
    "m3r_iter_proj_noalloc": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _f, _vp]),
    "m3r_refine_matches_f16": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "m3r_refine_matches_f32": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "m3r_match": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _f, _f, _i, _i, _vp]),
    "m3r_ray_image": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "m3r_constrain_points_to_ray": (_i, [_vp, _vp, _vp, _ll, _i, _i, _vp]),
    "m3r_update_pointmap": (_i, [_vp, _vp, _vp, _vp, _vp, _ll, _i, _vp]),
    "m3r_track_rays": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _ll, _d, _d, _d, _i, _d, _d, _vp, _vp, _vp, _vp, _vp]),
    "m3r_track_calib": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _ll, _i, _i, _d, _d, _d, _d, _d, _i, _d, _d,
                             _vp, _vp, _vp, _vp, _vp]),
    "m3r_gn_workspace_bytes": (_sz, [_i, _i, _i]),
    "m3r_gauss_newton": (_i, [_PP, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "m3r_gn_begin": (_i, [_PP, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _i, _vp, _sz, _vp]),
    "m3r_gn_accumulate": (_i, [_PP, _vp, _i, _i, _i, _vp, _vp]),
    "m3r_gn_iterate_peer": (_i, [_PP, _vp, _i, _i, _i, _vp, _i, _i, _vp, _vp, C.c_uint64, _vp]),
    "m3r_gn_hessian_pass": (_i, [_PP, _vp, _i, _i, _i, _vp, _vp]),
    "m3r_gn_system_ptr": (_vp, [_vp, _i, _i, _i]),
    "m3r_gn_system_count": (_sz, [_i]),
    "m3r_gn_record_count": (_ll, [_vp, _i, _i, _i]),
    "m3r_gn_solve_retract": (_i, [_PP, _vp, _i, _i, _i, _vp, _vp]),
    "m3r_gn_end": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp]),
    "m3r_gn_edge_blocks": (_i, [_PP, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "m3r_pose_retr": (_i, [_vp, _vp, _i, _i, _vp]),
    "m3r_iter_proj_host": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _f]),
    "m3r_refine_matches_f16_host": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i]),
    "m3r_gauss_newton_host": (_i, [_PP, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp]),
}

_lib = None


def load():
    """Load the shared library and bind every declared symbol. Raises if the library or a symbol is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python __graft_entry__.py build` "
            "(make -C mast3r_slam_ysh_b200/csrc). There is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale
        fn.restype = res
        fn.argtypes = args
    if lib.m3r_abi_version() != 1:
        raise ImportError("libm3r_backends_b200.so: ABI version mismatch")
    if os.environ.get("M3R_SOLVER") == "dense":  # A/B switch for benchmarking: envelope Cholesky off
        lib.m3r_debug_set_solver(1)
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().m3r_last_error_string().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")
