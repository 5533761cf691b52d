#!/usr/bin/env python
"""Builds the compiled binding mast3r_slam_ysh_b200/mast3r_slam_backends_b200*.so in-tree (g++ only; the kernels are in
libm3r_backends_b200.so, which it links with an $ORIGIN rpath).  Equivalent of the reference's setup.py:38-52 with the
gencode list gone (no device code here) and Eigen dropped (the Cholesky runs on the device)."""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(os.path.dirname(HERE))
ROOT = os.path.dirname(PKG)
NAME = "mast3r_slam_backends_b200"


def main():
    src = os.path.join(HERE, "gn_b200.cpp")
    out = os.path.join(PKG, NAME + sysconfig.get_config_var("EXT_SUFFIX"))
    lib = os.path.join(PKG, "libm3r_backends_b200.so")
    deps = [src, os.path.join(ROOT, "include", "m3r_backends.h"), __file__]
    if os.path.exists(out) and os.path.getmtime(out) > max(os.path.getmtime(d) for d in deps):
        print(f"[build_ext] up to date: {out}")
        return 0
    assert os.path.exists(lib), "build libm3r_backends_b200.so first (make -C mast3r_slam_ysh_b200/csrc)"
    from torch.utils import cpp_extension as ce
    inc = ce.include_paths("cuda") + [sysconfig.get_paths()["include"], os.path.join(ROOT, "include")]
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", src, "-o", out,
           f"-DTORCH_EXTENSION_NAME={NAME}", "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=1"]
    cmd += [f"-I{i}" for i in inc]
    cmd += [f"-L{p}" for p in ce.library_paths("cuda")] + [f"-L{PKG}", "-Wl,-rpath,$ORIGIN"]
    cmd += ["-lm3r_backends_b200", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", "-lcudart"]
    subprocess.check_call(cmd)
    print(f"[build_ext] built {out}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
