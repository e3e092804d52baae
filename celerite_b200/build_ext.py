"""In-tree build of the pybind11 module `celerite_b200.solver_ext` (host C++ class over the C ABI).

    python -m celerite_b200.build_ext

g++ only (no nvcc): the module links against libcelerite_b200.so next to it ($ORIGIN rpath).
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, "csrc", "host", "solver_ext.cpp")
HDR = os.path.join(HERE, "csrc", "host", "cholesky_solver.hpp")
OUT = os.path.join(HERE, "solver_ext" + sysconfig.get_config_var("EXT_SUFFIX"))
LIB = os.path.join(HERE, "libcelerite_b200.so")


def build(force=False, verbose=False):
    import pybind11
    if not os.path.exists(LIB):
        from . import build as _b
        _b.build()
    deps = [SRC, HDR, os.path.join(ROOT, "include", "celerite_b200.h"), LIB]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in deps):
        if verbose:
            print("up to date:", OUT)
        return OUT
    cxx = os.environ.get("CXX", "g++")
    cmd = [cxx, "-O2", "-shared", "-fPIC", "-std=c++17", "-fvisibility=hidden",
           "-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"],
           "-I", os.path.join(ROOT, "include"), SRC, "-o", OUT,
           "-L", HERE, "-lcelerite_b200", "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("solver_ext build failed:\n%s\n%s" % (r.stdout, r.stderr))
    if verbose:
        print("built", OUT)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
