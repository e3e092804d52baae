"""Drop-in check at the level a celerite user sees: the REFERENCE's own Python front-end
(celerite.GP, terms; installed unmodified into baseline/_ref) running on this library's solver
class instead of its pybind11 extension.  Mirrors tests/test_celerite.py of the reference:
log_likelihood / apply_inverse / dot / sample against dense linear algebra, LinAlgError and
`quiet`, recompute on dirty parameters, duplicate time stamps, pickling.

Both host layers are exercised: the pybind11 module built on the C++ host class
(celerite_b200.solver_ext) and the ctypes mirror (celerite_b200.solver)."""
import os
import pickle
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")


def _celerite():
    if not os.path.isdir(os.path.join(REF, "celerite")):
        pytest.skip("reference package not installed under baseline/_ref")
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import celerite
    return celerite


@pytest.fixture(params=["solver_ext", "solver"])
def cel(request, monkeypatch):
    celerite = _celerite()
    import importlib
    mod = importlib.import_module("celerite_b200." + request.param)
    monkeypatch.setattr(celerite.solver, "CholeskySolver", mod.CholeskySolver)
    monkeypatch.setattr(celerite.solver, "LinAlgError", mod.LinAlgError)
    return celerite


def _dense(celerite, kernel, t, yerr):
    K = kernel.get_value(t[:, None] - t[None, :])
    K[np.diag_indices_from(K)] += yerr ** 2
    return K


def _kernels(terms):
    return [
        terms.RealTerm(log_a=0.1, log_c=0.5),
        terms.RealTerm(log_a=0.1, log_c=0.5) + terms.RealTerm(log_a=-0.1, log_c=0.7),
        terms.ComplexTerm(log_a=0.1, log_c=0.5, log_d=0.1),
        terms.ComplexTerm(log_a=0.1, log_b=-0.2, log_c=0.5, log_d=0.1),
        terms.SHOTerm(log_S0=0.1, log_Q=1.0, log_omega0=0.5),
        terms.SHOTerm(log_S0=0.1, log_Q=-1.0, log_omega0=0.5),        # Q < 1/2: two real terms
        terms.SHOTerm(log_S0=0.1, log_Q=1.0, log_omega0=0.5) + terms.RealTerm(log_a=0.1, log_c=0.4),
        terms.SHOTerm(log_S0=0.1, log_Q=1.0, log_omega0=0.5) * terms.RealTerm(log_a=0.1, log_c=0.4),
        terms.Matern32Term(log_sigma=0.1, log_rho=0.4),
    ]


def test_gp_log_likelihood_matches_dense(cel):
    rng = np.random.default_rng(42)
    t = np.sort(rng.uniform(0, 5, 100))
    yerr = rng.uniform(0.1, 0.5, len(t))
    y = np.sin(t)
    for kernel in _kernels(cel.terms):
        gp = cel.GP(kernel)
        gp.compute(t, yerr)
        K = _dense(cel, kernel, t, yerr)
        ll0 = -0.5 * (y @ np.linalg.solve(K, y) + np.linalg.slogdet(K)[1] + len(t) * np.log(2 * np.pi))
        assert np.allclose(gp.log_likelihood(y), ll0)
        assert np.allclose(gp.solver.log_determinant(), np.linalg.slogdet(K)[1])
        # dirty-flag recompute (tests/test_celerite.py:372-391)
        v = gp.get_parameter_vector()
        gp.set_parameter_vector(v + 0.1)
        K2 = _dense(cel, gp.kernel, t, yerr)
        ll2 = -0.5 * (y @ np.linalg.solve(K2, y) + np.linalg.slogdet(K2)[1] + len(t) * np.log(2 * np.pi))
        assert np.allclose(gp.log_likelihood(y), ll2)


def test_gp_apply_inverse_dot_sample(cel):
    rng = np.random.default_rng(1)
    t = np.sort(rng.uniform(0, 5, 300))
    yerr = rng.uniform(0.1, 0.5, len(t))
    kernel = cel.terms.SHOTerm(log_S0=0.1, log_Q=1.0, log_omega0=0.5) + cel.terms.RealTerm(log_a=0.1, log_c=0.4)
    gp = cel.GP(kernel)
    gp.compute(t, yerr)
    K = _dense(cel, kernel, t, yerr)
    b = rng.standard_normal((len(t), 4))
    assert np.allclose(gp.apply_inverse(b), np.linalg.solve(K, b))
    # solve() always returns N x nrhs (cholesky.h:218-223), also through GP.apply_inverse
    assert np.allclose(gp.apply_inverse(b[:, 0])[:, 0], np.linalg.solve(K, b[:, 0]))
    K0 = kernel.get_value(t[:, None] - t[None, :])
    assert np.allclose(gp.dot(b, t), K0 @ b)                         # tests/test_celerite.py:154-192
    s = gp.sample(size=3)                                            # uses solver.dot_L
    assert s.shape == (3, len(t)) and np.all(np.isfinite(s))


def test_gp_predict(cel):
    """tests/test_celerite.py:468-496: GP.predict mean (O(N+M) solver.predict path) vs dense."""
    rng = np.random.default_rng(11)
    t = np.sort(rng.uniform(0, 5, 200))
    yerr = rng.uniform(0.1, 0.5, len(t))
    y = np.sin(t)
    kernel = cel.terms.RealTerm(0.1, 0.5) + cel.terms.ComplexTerm(0.6, 0.7, 1.0)
    gp = cel.GP(kernel)
    gp.compute(t, yerr)
    x = np.linspace(-1.0, 6.0, 500)
    K = _dense(cel, kernel, t, yerr)
    Ks = kernel.get_value(x[:, None] - t[None, :])
    mu0 = Ks @ np.linalg.solve(K, y)
    mu = gp.predict(y, x, return_cov=False, return_var=False)
    assert np.allclose(mu, mu0)


def test_gp_not_positive_definite_and_quiet(cel):
    """tests/test_celerite.py:326-345: LinAlgError from compute and log_likelihood, -inf when
    quiet, RuntimeError before compute."""
    rng = np.random.default_rng(3)
    x = np.sort(rng.uniform(0, 1, 10))
    y = np.sin(x)

    class NPDTerm(cel.terms.Term):
        parameter_names = ("par1", )

        def get_real_coefficients(self, params):  # NOQA
            return [params[0]], [0.1]

    gp = cel.GP(NPDTerm(-1.0))
    with pytest.raises(cel.solver.LinAlgError):
        gp.compute(x, 0.0)
    with pytest.raises(cel.solver.LinAlgError):
        gp.log_likelihood(y)
    assert np.isinf(gp.log_likelihood(y, quiet=True))

    gp = cel.GP(cel.terms.RealTerm(0.1, 0.5))
    with pytest.raises(RuntimeError):
        gp.log_likelihood(y)


def test_gp_zero_dt(cel):
    """tests/test_celerite.py:393-404: a duplicated time stamp."""
    rng = np.random.default_rng(5)
    t = np.sort(rng.uniform(0, 5, 60))
    t = np.sort(np.append(t, t[10]))
    yerr = 0.3 * np.ones_like(t)
    y = np.sin(t)
    kernel = cel.terms.SHOTerm(log_S0=0.1, log_Q=1.0, log_omega0=0.5)
    gp = cel.GP(kernel)
    gp.compute(t, yerr)
    K = _dense(cel, kernel, t, yerr)
    ll0 = -0.5 * (y @ np.linalg.solve(K, y) + np.linalg.slogdet(K)[1] + len(t) * np.log(2 * np.pi))
    assert np.allclose(gp.log_likelihood(y), ll0)


def test_gp_pickle(cel):
    """tests/test_celerite.py:237-289"""
    rng = np.random.default_rng(7)
    t = np.sort(rng.uniform(0, 5, 80))
    y = np.sin(t)
    gp = cel.GP(cel.terms.RealTerm(log_a=0.1, log_c=0.5))
    gp2 = pickle.loads(pickle.dumps(gp, -1))
    assert not gp2.computed
    gp.compute(t, 0.2)
    ll = gp.log_likelihood(y)
    gp3 = pickle.loads(pickle.dumps(gp, -1))
    assert gp3.computed and np.allclose(gp3.log_likelihood(y), ll)


def test_solver_level_gradient_matches_reference_extension(cel):
    """The reference's own CPU extension (forward-mode AD) against ours on the same inputs."""
    celerite = cel
    import importlib
    import importlib.util
    from celerite_b200 import solver_ext
    rng = np.random.default_rng(9)
    N = 400
    t = np.sort(rng.uniform(0, 20, N)); yerr = rng.uniform(0.1, 0.4, N); y = rng.standard_normal(N)
    kernel = celerite.terms.SHOTerm(log_S0=0.1, log_Q=1.0, log_omega0=0.5) + celerite.terms.RealTerm(log_a=0.1, log_c=0.4)
    coeffs = kernel.coefficients
    E0, E2 = np.empty(0), np.empty((0, 0))
    ours = solver_ext.CholeskySolver()
    v1, g1 = ours.grad_log_likelihood(0.01, *coeffs, E0, E2, E2, t, y, yerr ** 2)
    # the fixture patched celerite.solver.CholeskySolver: reach the reference implementation
    # through a fresh load of the compiled module
    spec = importlib.util.find_spec("celerite.solver")
    fresh = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fresh)
    v0, g0 = fresh.CholeskySolver().grad_log_likelihood(0.01, *coeffs, E0, E2, E2, t, y, yerr ** 2)
    assert abs(v1 - v0) / abs(v0) < 1e-10
    assert np.max(np.abs(g1 - g0) / np.maximum(np.abs(g0), 1e-12 * np.abs(g0).max())) < 1e-8
