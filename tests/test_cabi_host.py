"""CPU tests of the boundary: the C-ABI library loads and exports every symbol include/m3r_backends.h declares, argument
validation works without a GPU (no compute calls), and the host mirror rejects what the reference rejects."""
import ctypes as C
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "m3r_backends.h")).read()
    return sorted(set(re.findall(r"^M3R_API [\w\* ]+?(m3r_\w+)\(", text, flags=re.M)))


def test_header_declares_expected_entry_points():
    names = _declared()
    for must in ("m3r_iter_proj", "m3r_refine_matches_f16", "m3r_gauss_newton", "m3r_gn_begin", "m3r_gn_accumulate",
                 "m3r_gn_solve_retract", "m3r_gn_end", "m3r_gn_edge_blocks", "m3r_pose_retr", "m3r_gauss_newton_host"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from mast3r_slam_ysh_b200 import _lib
    lib = C.CDLL(_lib.LIB_PATH)
    for name in _declared():
        assert hasattr(lib, name), name
    assert set(_lib.SYMBOLS) == set(_declared())
    assert _lib.load().m3r_abi_version() == 1


def test_argument_validation_without_gpu():
    from mast3r_slam_ysh_b200 import _lib
    lib = _lib.load()
    # empty problems are no-ops; bad sizes are rejected before any CUDA call
    assert lib.m3r_iter_proj(None, None, None, None, None, 0, 384, 512, 0, 10, 1e-8, 1e-6, None) == 0
    assert lib.m3r_iter_proj(None, None, None, None, None, 1, 384, 512, 16, 10, 1e-8, 1e-6, None) == -1
    assert b"null" in lib.m3r_last_error_string()
    assert lib.m3r_refine_matches_f16(None, None, None, None, 0, 4, 4, 0, 24, 3, 5, None) == 0
    assert lib.m3r_refine_matches_f16(None, None, None, None, 1, 4, 4, 4, 24, 3, 5, None) == -1
    p = _lib.GnParams(kind=7)
    assert lib.m3r_gauss_newton(C.byref(p), None, None, None, None, None, None, None, None, 2, 4, 1, None, None, None, 0, None) == -1
    assert b"unknown kind" in lib.m3r_last_error_string()
    p = _lib.GnParams(kind=_lib.M3R_GN_RAYS, max_iter=1)
    buf = (C.c_char * 4096)()
    addr = (C.addressof(buf) + 255) // 256 * 256
    rc = lib.m3r_gauss_newton(C.byref(p), C.c_void_p(addr), C.c_void_p(addr), C.c_void_p(addr), C.c_void_p(addr),
                              C.c_void_p(addr), C.c_void_p(addr), C.c_void_p(addr), C.c_void_p(addr), 2, 4, 1, None,
                              None, C.c_void_p(addr), 16, None)
    assert rc == -2 and b"workspace too small" in lib.m3r_last_error_string()


def test_workspace_size_is_monotone():
    from mast3r_slam_ysh_b200 import _lib
    lib = _lib.load()
    a = lib.m3r_gn_workspace_bytes(10, 196608, 48)
    b = lib.m3r_gn_workspace_bytes(32, 196608, 512)
    assert 0 < a < b
    assert b < 1.01 * (512 * 196608 * 32) + (1 << 24)  # dominated by the 32-byte edge records (capacity E x N)
    assert lib.m3r_gn_system_count(32) >= (7 * 31 + 1) * 7 * 31


def test_host_mirror_rejects_cpu_and_noncontiguous_tensors():
    from mast3r_slam_ysh_b200 import backends as be
    rays = torch.zeros(1, 8, 8, 9)
    with pytest.raises(RuntimeError, match="CUDA"):
        be.iter_proj(rays, torch.zeros(1, 64, 3), torch.zeros(1, 64, 2), 10, 1e-8, 1e-6)
    with pytest.raises(RuntimeError, match="contiguous"):  # CHECK_CONTIGUOUS (gn.h:5)
        be.iter_proj(rays.permute(0, 2, 1, 3), torch.zeros(1, 64, 3), torch.zeros(1, 64, 2), 10, 1e-8, 1e-6)
    with pytest.raises(RuntimeError, match="contiguous"):
        be.refine_matches(torch.zeros(1, 8, 8, 24).half().transpose(1, 2), torch.zeros(1, 64, 24).half(),
                          torch.zeros(1, 64, 2, dtype=torch.int64), 3, 5)


def test_drop_in_module_name():
    import mast3r_slam_backends as m  # the name the reference imports (matching.py:5, global_opt.py:9)
    for fn in ("gauss_newton_points", "gauss_newton_rays", "gauss_newton_calib", "iter_proj", "refine_matches"):
        assert callable(getattr(m, fn))  # gn.cpp:116-123


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "mast3r_slam_ysh_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("# oracle", ""), f


def test_compiled_binding_exports_the_reference_function_table():
    """The compiled torch extension (csrc/ext/gn_b200.cpp) is the pybind module of gn.cpp:116-123: same five names, and
    like the reference's CHECK_CONTIGUOUS (gn.h:5) it raises RuntimeError -- also for CPU tensors (no CPU path)."""
    import importlib
    mod = importlib.import_module("mast3r_slam_ysh_b200.mast3r_slam_backends_b200")
    for name in ("gauss_newton_points", "gauss_newton_rays", "gauss_newton_calib", "iter_proj", "refine_matches"):
        assert callable(getattr(mod, name))
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        mod.iter_proj(torch.zeros(1, 4, 4, 9), torch.zeros(1, 2, 3), torch.zeros(1, 2, 2), 10, 1e-8, 1e-6)
    with pytest.raises(RuntimeError, match="contiguous"):
        mod.iter_proj(torch.zeros(1, 4, 9, 4).permute(0, 1, 3, 2), torch.zeros(1, 2, 3), torch.zeros(1, 2, 2), 10, 1e-8, 1e-6)
    with pytest.raises(TypeError):  # positional signature of gn.h:112-117
        mod.refine_matches(torch.zeros(1), torch.zeros(1))
    import sys
    sys.path.insert(0, ROOT)
    import mast3r_slam_backends as shim
    assert shim.BINDING in ("compiled", "ctypes")
    if os.environ.get("M3R_BINDING", "auto") == "auto":
        assert shim.BINDING == "compiled" and shim.iter_proj is mod.iter_proj
