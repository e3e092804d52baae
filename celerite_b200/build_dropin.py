"""Builds the reference's OWN sources against the drop-in header (csrc/host/dropin).

    python -m celerite_b200.build_dropin [--force]

Nothing of the reference is copied into the repository: its files are compiled where they lie
(/root/reference) and every output goes to baseline/_dropin/, which is git-ignored but travels to
the GPU box.  What is built:

  baseline/_dropin/celerite/            the reference's Python package, unmodified, whose extension
      solver.*.so                       module is celerite/solver.cpp compiled UNCHANGED with
                                        celerite_b200/csrc/host/dropin in front of cpp/include:
                                        `celerite.solver.CholeskySolver` then runs on the GPU while
                                        get_kernel_value, get_psd_value, check_coefficients and
                                        CARMASolver stay the reference's own code (solver.cpp:89-235)
  baseline/_dropin/ref_tests/           the reference's tests/test_celerite.py, tests/test_terms.py
  baseline/_dropin/bin/test_solvers     cpp/src/test_solvers.cc, unchanged (Cholesky vs DirectSolver)
  baseline/_dropin/bin/test_grad        cpp/src/test_grad.cc, unchanged, and test_grad_ref = the same
                                        file against the reference's header alone (its printed
                                        derivatives are what ours must reproduce)
  baseline/_dropin/bin/benchmark        cpp/src/benchmark.cc, unchanged, and benchmark_ref
  baseline/_dropin/bin/test_overloads   tests/cpp/test_dropin_overloads.cc (ours): the five compute
                                        overloads of solver.h:96-207 and every other method against
                                        the reference's DirectSolver / CholeskySolver on the CPU
"""
import glob
import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("CELERITE_REF", "/root/reference")
OUT = os.path.join(ROOT, "baseline", "_dropin")
SHADOW = os.path.join(HERE, "csrc", "host", "dropin")
LIB = os.path.join(HERE, "libcelerite_b200.so")
CXX = os.environ.get("CXX", "g++")


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("%s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))


def _newer(out, deps):
    return os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps)


def build(force=False, verbose=False):
    if not os.path.isdir(os.path.join(REF, "cpp", "include")):
        if verbose:
            print("reference tree not mounted at %s: keeping prebuilt baseline/_dropin (if any)" % REF)
        return None
    import pybind11
    if not os.path.exists(LIB):
        from . import build as _b
        _b.build()
    inc_ref = ["-I", os.path.join(REF, "cpp", "include"), "-I", os.path.join(REF, "cpp", "lib", "eigen")]
    inc_ours = ["-I", SHADOW, "-I", os.path.join(ROOT, "include")] + inc_ref
    shadow_hdr = os.path.join(SHADOW, "celerite", "solver", "cholesky.h")
    abi_hdr = os.path.join(ROOT, "include", "celerite_b200.h")
    link = ["-L", HERE, "-lcelerite_b200"]
    pkg = os.path.join(OUT, "celerite")
    bindir = os.path.join(OUT, "bin")
    os.makedirs(pkg, exist_ok=True)
    os.makedirs(bindir, exist_ok=True)
    os.makedirs(os.path.join(OUT, "ref_tests"), exist_ok=True)

    # the reference's Python package and tests, as they are
    for f in glob.glob(os.path.join(REF, "celerite", "*.py")):
        shutil.copy(f, pkg)
    with open(os.path.join(pkg, "celerite_version.py"), "w") as f:   # setuptools_scm would write this
        f.write('__version__ = "0.4.3.dev0+reference.b200"\n')
    for f in ("test_celerite.py", "test_terms.py"):
        shutil.copy(os.path.join(REF, "tests", f), os.path.join(OUT, "ref_tests"))

    jobs = []
    ext = os.path.join(pkg, "solver" + sysconfig.get_config_var("EXT_SUFFIX"))
    src = os.path.join(REF, "celerite", "solver.cpp")
    jobs.append((ext, [src, shadow_hdr, abi_hdr, LIB],
                 [CXX, "-O2", "-DNDEBUG", "-shared", "-fPIC", "-std=c++14", "-fvisibility=hidden",   # NDEBUG as in a stock setup.py build
                  "-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"]] + inc_ours +
                 [src, "-o", ext] + link + ["-Wl,-rpath,$ORIGIN/../../../celerite_b200"]))
    rpath_bin = ["-Wl,-rpath,$ORIGIN/../../../celerite_b200"]
    for name in ("test_solvers", "test_grad", "benchmark"):
        src = os.path.join(REF, "cpp", "src", name + ".cc")
        out = os.path.join(bindir, name)
        jobs.append((out, [src, shadow_hdr, abi_hdr, LIB],
                     [CXX, "-O2", "-std=c++14"] + inc_ours + [src, "-o", out] + link + rpath_bin))
        if name != "test_solvers":
            out = os.path.join(bindir, name + "_ref")
            jobs.append((out, [src], [CXX, "-O2", "-march=x86-64-v3", "-DNDEBUG", "-std=c++14"] + inc_ref + [src, "-o", out]))
    src = os.path.join(ROOT, "tests", "cpp", "test_dropin_overloads.cc")
    out = os.path.join(bindir, "test_overloads")
    jobs.append((out, [src, shadow_hdr, abi_hdr, LIB],
                 [CXX, "-O2", "-std=c++14"] + inc_ours + [src, "-o", out] + link + rpath_bin))

    import concurrent.futures
    todo = [(o, c) for o, d, c in jobs if force or not _newer(o, d)]
    with concurrent.futures.ThreadPoolExecutor(max_workers=4) as ex:
        list(ex.map(lambda oc: _run(oc[1]), todo))
    if verbose:
        print("drop-in build: %d of %d targets rebuilt under %s" % (len(todo), len(jobs), OUT))
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
