"""The reference's OWN code on top of the drop-in header (celerite_b200/csrc/host/dropin).

celerite_b200/build_dropin.py compiles, unchanged and where they lie, the reference's pybind11
extension (celerite/solver.cpp) and its C++ programs (cpp/src/test_solvers.cc, test_grad.cc,
benchmark.cc) with the drop-in directory in front of cpp/include, so that
`celerite::solver::CholeskySolver<double>` (and the AutoDiffScalar instantiation the gradient
lambda uses) is the CUDA implementation.  The outputs live in baseline/_dropin (git-ignored, built
in the container, shipped to the GPU box).  Here they run on the GPU:

  * the reference's unmodified tests/test_celerite.py (every with_general=True case included) and
    tests/test_terms.py against `celerite.solver` = that extension -- no sys.modules patching;
  * one script executed twice, once per install (baseline/_ref = stock CPU reference,
    baseline/_dropin = same Python, GPU solver), outputs compared at north_star's tolerances;
  * the C++ programs: test_solvers (Cholesky vs the dense DirectSolver, 1e-10), our
    tests/cpp/test_dropin_overloads.cc (five compute overloads of solver.h:96-207 etc.), and
    test_grad, whose printed AutoDiff derivatives must equal those of the same file built against
    the reference's header alone.
"""
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "baseline", "_dropin")
REFPKG = os.path.join(ROOT, "baseline", "_ref")


def _need(*paths):
    for p in paths:
        if not os.path.exists(p):
            pytest.skip("%s is not built (python -m celerite_b200.build_dropin needs /root/reference)" % p)


def _env(pkg):
    env = dict(os.environ)
    env["PYTHONPATH"] = pkg
    env.pop("PYTEST_ADDOPTS", None)
    return env


def _run_suite(test_file, pkg):
    _need(os.path.join(pkg, "celerite", "__init__.py"), os.path.join(DROPIN, "ref_tests", test_file))
    r = subprocess.run([sys.executable, "-m", "pytest", test_file, "-q", "-p", "no:cacheprovider",
                        "-o", "addopts=", "--rootdir", os.path.join(DROPIN, "ref_tests")],
                       cwd=os.path.join(DROPIN, "ref_tests"), env=_env(pkg), capture_output=True, text=True,
                       timeout=900)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    m = re.search(r"(\d+) passed", r.stdout)
    assert m, tail
    assert "failed" not in r.stdout.splitlines()[-1], tail
    return int(m.group(1))


@pytest.mark.parametrize("test_file", ["test_celerite.py", "test_terms.py"])
def test_reference_pytest_file_unmodified_on_the_gpu_solver(test_file):
    """Same file, same interpreter, two installs: every test the stock CPU install passes must pass on
    the GPU-backed `celerite.solver` (35 in tests/test_celerite.py with this pytest: its
    itertools.product parametrisation of test_grad_log_likelihood is consumed at collection)."""
    n_gpu = _run_suite(test_file, DROPIN)
    n_cpu = _run_suite(test_file, REFPKG)
    assert n_gpu == n_cpu and n_gpu >= 1


SCRIPT = r"""
import json, sys, pickle
import numpy as np
import celerite
from celerite import terms, solver
out = {"pkg": celerite.__file__}
rng = np.random.RandomState(7)
N = 700
t = np.sort(rng.uniform(0, 20, N)); yerr = rng.uniform(0.1, 0.3, N); y = np.sin(t) + yerr * rng.randn(N)
kernel = terms.RealTerm(log_a=0.1, log_c=-0.5) + terms.ComplexTerm(log_a=0.3, log_b=-1.0, log_c=-0.7, log_d=0.4) \
    + terms.SHOTerm(log_S0=0.2, log_Q=1.1, log_omega0=0.9) + terms.JitterTerm(log_sigma=-2.0)
gp = celerite.GP(kernel, mean=0.25, fit_mean=True)
gp.compute(t, yerr)
out["ll"] = float(gp.log_likelihood(y))
out["apply_inverse"] = gp.apply_inverse(y).ravel().tolist()
xs = np.linspace(-1, 21, 333)
mu, var = gp.predict(y, xs, return_var=True)
out["predict_mu"] = mu.tolist(); out["predict_var"] = var.tolist()
out["dot"] = gp.dot(y).ravel().tolist()
np.random.seed(3)
out["sample"] = gp.sample().tolist()
# raw solver: gradient with general rows (the with_general=True shape of the reference's tests)
U = np.vander(t / 20.0, 3).T.copy(); V = U * 0.5; A = 1e-6 + np.sum(U * V, axis=0)
c = kernel.coefficients
s = solver.CholeskySolver()
val, grad = s.grad_log_likelihood(kernel.jitter, c[0], c[1], c[2], c[3], c[4], c[5], A, U, V, t, y, yerr ** 2)
out["grad_val"] = float(val); out["grad"] = grad.tolist()
s.compute(kernel.jitter, c[0], c[1], c[2], c[3], c[4], c[5], A, U, V, t, yerr ** 2)
out["general_logdet"] = float(s.log_determinant()); out["general_dot_solve"] = float(s.dot_solve(y))
out["dot_L"] = s.dot_L(y).ravel().tolist()
# pickle round trip through the reference's own serialize/deserialize
s2 = pickle.loads(pickle.dumps(s, -1))
out["pickled_dot_solve"] = float(s2.dot_solve(y))
state = s.__getstate__()
out["state_shapes"] = [list(np.shape(a)) for a in state[4:]]
out["psd"] = float(solver.get_psd_value(c[0], c[1], c[2], c[3], c[4], c[5], np.array([0.3]))[0])
out["has_carma"] = hasattr(solver, "CARMASolver")
out["maps"] = sorted({l.split("/")[-1].strip() for l in open("/proc/self/maps") if "celerite" in l})
print("RESULT" + json.dumps(out))
"""


def _run_script(pkg):
    r = subprocess.run([sys.executable, "-c", SCRIPT], env=_env(pkg), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT")][-1]
    return json.loads(line[len("RESULT"):])


def test_same_python_two_installs_gpu_against_cpu_reference():
    _need(os.path.join(DROPIN, "celerite", "__init__.py"), os.path.join(REFPKG, "celerite", "__init__.py"))
    ours, ref = _run_script(DROPIN), _run_script(REFPKG)
    assert ours["pkg"].startswith(DROPIN) and ref["pkg"].startswith(REFPKG)
    assert "libcelerite_b200.so" in ours["maps"], ours["maps"]          # the CUDA library really is underneath
    assert "libcelerite_b200.so" not in ref["maps"]
    assert ours["has_carma"] and ours["state_shapes"] == ref["state_shapes"]
    for key in ("ll", "grad_val", "general_logdet", "general_dot_solve", "pickled_dot_solve", "psd"):
        assert abs(ours[key] - ref[key]) <= 1e-10 * abs(ref[key]), (key, ours[key], ref[key])
    for key in ("apply_inverse", "predict_mu", "predict_var", "dot", "sample", "dot_L"):
        a, b = np.array(ours[key]), np.array(ref[key])
        assert np.max(np.abs(a - b)) <= 1e-10 * np.max(np.abs(b)), key
    g, gr = np.array(ours["grad"]), np.array(ref["grad"])
    assert np.max(np.abs(g - gr) / np.maximum(np.abs(gr), 1e-12 * np.max(np.abs(gr)))) < 1e-8


def _bin(name):
    p = os.path.join(DROPIN, "bin", name)
    _need(p)
    return p


@pytest.mark.parametrize("name", ["test_solvers", "test_overloads"])
def test_cpp_program_on_the_gpu_solver(name):
    r = subprocess.run([_bin(name)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert "FAILED" not in r.stderr and "Test failed" not in r.stderr
    if name == "test_solvers":      # the eight checks of cpp/src/test_solvers.cc:79-97
        assert r.stderr.count("Test passed") == 8, r.stderr


def test_cpp_autodiff_program_prints_the_reference_derivatives():
    """cpp/src/test_grad.cc prints d logdet / d theta and d (y^T K^-1 y) / d theta for three kernels (the
    second and third with general rows) and asserts nothing itself; built against the reference's header
    alone (test_grad_ref, forward-mode AD on the CPU) the same file gives the numbers to match."""
    ours = subprocess.run([_bin("test_grad")], capture_output=True, text=True, timeout=600)
    ref = subprocess.run([_bin("test_grad_ref")], capture_output=True, text=True, timeout=600)
    assert ours.returncode == 0 and ref.returncode == 0, ours.stderr[-2000:]
    a = np.array([float(v) for v in ours.stdout.split()])
    b = np.array([float(v) for v in ref.stdout.split()])
    assert a.shape == b.shape and a.size == 6 * 9
    # std::cout prints 6 significant digits; compare at that precision, block by block
    for blk_a, blk_b in zip(a.reshape(6, 9), b.reshape(6, 9)):
        assert np.max(np.abs(blk_a - blk_b)) <= 2e-5 * np.max(np.abs(blk_b)), (blk_a, blk_b)


def test_reference_benchmark_program_runs_on_the_gpu_solver():
    r = subprocess.run([_bin("benchmark"), "3", "4096", "2"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-2000:]
    assert len(r.stdout.strip().splitlines()) >= 7        # one line per N = 64 ... 4096
