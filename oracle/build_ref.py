#!/usr/bin/env python
"""TEST INFRASTRUCTURE ONLY -- builds the *reference's own* CUDA backend as `mast3r_slam_backends_ref`.

The reference ships only sm_60..sm_86 gencodes and uses two torch C++ APIs that are gone in torch 2.11
(SURVEY.md section 8c).  This recipe
  1. copies `mast3r_slam/backend/{src,include}` from /root/reference into a temp dir (never into this repo),
  2. applies the 4-line API patch (`D11.type()` -> `D11.scalar_type()`, matching_kernels.cu:103;
     `torch::linalg::linalg_norm` -> `at::linalg_norm`, gn_kernels.cu:802,1219,1629),
  3. compiles it for sm_100a with nvcc/g++ directly (not the reference's setup.py),
  4. writes ONLY into `oracle/_ref/` (git-ignored; it travels to the GPU box): `mast3r_slam_backends_ref*.so` and
     `pyref/` = unchanged copies of the reference's pure-Python callers of the path (see stage_python).

The resulting module is the GPU oracle for `tests/ -m gpu` and the "reference CUDA on the same B200" timing arm
in bench.py.  Nothing in the product package imports it.  It cannot run in a GPU-less container.
"""
import os
import shutil
import subprocess
import sys
import sysconfig
import tempfile

REF = os.environ.get("M3R_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
NAME = "mast3r_slam_backends_ref"


def patch(text, old, new, count):
    assert text.count(old) == count, (old, text.count(old), count)
    return text.replace(old, new)


PY_STAGE = ("mast3r_slam/__init__.py", "mast3r_slam/matching.py", "mast3r_slam/image.py", "mast3r_slam/config.py",
            "mast3r_slam/nonlinear_optimizer.py", "config/base.yaml")


def stage_python():
    """The reference's pure-Python callers of the path (matching.py and what it imports, nonlinear_optimizer.py,
    config/base.yaml) are staged UNCHANGED under oracle/_ref/pyref/ (git-ignored like the .so, travels to the GPU box)
    so that tests can run them over the drop-in shim where /root/reference does not exist."""
    dst_root = os.path.join(OUT, "pyref")
    for rel in PY_STAGE:
        src = os.path.join(REF, rel)
        dst = os.path.join(dst_root, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            shutil.copyfile(src, dst)
    print(f"[build_ref] staged {len(PY_STAGE)} reference Python files under {dst_root}")


def main():
    src = os.path.join(REF, "mast3r_slam", "backend")
    if not os.path.isdir(src):
        print(f"[build_ref] {src} not present; skipping (prebuilt oracle/_ref is used if it exists)")
        return 0
    os.makedirs(OUT, exist_ok=True)
    stage_python()
    so = os.path.join(OUT, NAME + sysconfig.get_config_var("EXT_SUFFIX"))
    newest_src = max(
        os.path.getmtime(os.path.join(src, "src", f)) for f in ("gn.cpp", "gn_kernels.cu", "matching_kernels.cu")
    )
    if os.path.exists(so) and os.path.getmtime(so) > max(newest_src, os.path.getmtime(__file__)):
        print(f"[build_ref] up to date: {so}")
        return 0

    import torch  # noqa: F401  (only for include/library paths)
    from torch.utils import cpp_extension as ce

    tmp = tempfile.mkdtemp(prefix="m3r_ref_build_")
    try:
        shutil.copytree(os.path.join(src, "src"), os.path.join(tmp, "src"))
        shutil.copytree(os.path.join(src, "include"), os.path.join(tmp, "include"))
        p = os.path.join(tmp, "src", "matching_kernels.cu")
        t = open(p).read()
        t = patch(t, "AT_DISPATCH_FLOATING_TYPES_AND_HALF(D11.type(),", "AT_DISPATCH_FLOATING_TYPES_AND_HALF(D11.scalar_type(),", 1)
        open(p, "w").write(t)
        p = os.path.join(tmp, "src", "gn_kernels.cu")
        t = open(p).read()
        t = patch(t, "torch::linalg::linalg_norm(", "at::linalg_norm(", 3)
        open(p, "w").write(t)

        inc = ce.include_paths("cuda") + [
            sysconfig.get_paths()["include"],
            os.path.join(tmp, "include"),
            os.path.join(REF, "thirdparty", "eigen"),
        ]
        incf = [f"-I{i}" for i in inc]
        defs = [f"-DTORCH_EXTENSION_NAME={NAME}", "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=1"]
        nvcc_flags = [
            "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
            "-D__CUDA_NO_HALF_OPERATORS__", "-D__CUDA_NO_HALF_CONVERSIONS__",
            "-D__CUDA_NO_BFLOAT16_CONVERSIONS__", "-D__CUDA_NO_HALF2_OPERATORS__",
            "-gencode", "arch=compute_100a,code=sm_100a", "-w",
        ]
        objs = []
        procs = []
        for f in ("gn_kernels.cu", "matching_kernels.cu"):
            o = os.path.join(tmp, f + ".o")
            objs.append(o)
            cmd = ["nvcc", "-c", os.path.join(tmp, "src", f), "-o", o] + nvcc_flags + defs + incf
            procs.append(subprocess.Popen(cmd))
        o = os.path.join(tmp, "gn.cpp.o")
        objs.append(o)
        cmd = ["g++", "-c", os.path.join(tmp, "src", "gn.cpp"), "-o", o, "-O2", "-std=c++17", "-fPIC", "-w"] + defs + incf
        procs.append(subprocess.Popen(cmd))
        for pr in procs:
            if pr.wait() != 0:
                raise RuntimeError("reference compile failed")
        libs = [f"-L{p}" for p in ce.library_paths("cuda")]
        cmd = ["g++", "-shared", "-o", so] + objs + libs + [
            "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", "-lcudart",
        ]
        subprocess.check_call(cmd)
        print(f"[build_ref] built {so}")
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
