"""GPU parity, single-series API: CUDA CholeskySolver (through the C ABI) vs the reference's
outputs — the committed golden vectors and, where present, oracle/_ref or the C port on freshly
seeded inputs.  Tolerances are north_star's: 1e-10 relative on log-det / log-lik / solves."""
import pickle

import numpy as np
import pytest

import oracle
from tests.util import RTOL_SCALAR, RTOL_SOLVE, golden_names, rel_inf, rel_sum

pytestmark = pytest.mark.gpu
NAMES = golden_names()


@pytest.fixture(scope="module", params=["solver", "solver_ext"])
def S(request):
    """both host layers over the C ABI: the ctypes mirror and the pybind11/C++ host class"""
    import importlib
    return importlib.import_module("celerite_b200." + request.param)


def _compute(solver_mod, p):
    s = solver_mod.CholeskySolver()
    s.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp,
              p.A if p.A is not None else np.empty(0),
              p.U if p.U is not None else np.empty((0, 0)),
              p.V if p.V is not None else np.empty((0, 0)), p.x, p.diag)
    return s


@pytest.mark.parametrize("name", NAMES)
def test_golden(golden, S, name):
    c = golden[name]
    p = c.problem
    if c.status != 0:
        with pytest.raises(S.LinAlgError):
            _compute(S, p)
        return
    s = _compute(S, p)
    assert s.computed()
    ld = s.log_determinant()
    assert rel_sum(ld, c.log_det, p.N) < RTOL_SCALAR
    comp, N, J, ld2, phi, u, W, D = s.__getstate__()
    assert comp and N == p.N and J == p.J and ld2 == ld
    assert phi.shape == (p.J, max(p.N - 1, 0)) and W.shape == (p.J, p.N)
    if p.J and p.N > 1:
        assert rel_inf(phi.T, c.phi) < RTOL_SOLVE
        assert rel_inf(u.T, c.u) < RTOL_SOLVE
    if p.J:
        assert rel_inf(W.T, c.W) < RTOL_SOLVE
    assert rel_inf(D, c.D) < RTOL_SOLVE
    assert rel_sum(s.dot_solve(c.y), c.dot_solve, p.N) < RTOL_SCALAR
    z = s.solve(c.rhs)
    assert z.shape == c.solve.shape and z.flags.f_contiguous
    assert rel_inf(z, c.solve) < RTOL_SOLVE
    z1 = s.solve(c.y)
    assert z1.shape == (p.N, 1)  # always 2-D (cholesky.h:218-223)
    assert rel_inf(z1, c.solve_y) < RTOL_SOLVE
    assert rel_inf(s.dot_L(c.rhs), c.dot_L) < RTOL_SOLVE
    out = s.dot(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp,
                p.A if p.A is not None else np.empty(0),
                p.U if p.U is not None else np.empty((0, 0)),
                p.V if p.V is not None else np.empty((0, 0)), p.x, c.rhs)
    assert rel_inf(out, c.dot) < RTOL_SOLVE
    pred = s.predict(c.y, c.xs)                       # cholesky.h:599-698
    assert pred.shape == c.predict.shape
    assert rel_inf(pred, c.predict) < RTOL_SOLVE


def test_predict_needs_coefficients_after_unpickle(golden, S):
    """The pickled state holds no coefficients (solver.cpp:36-42): predict on an unpickled solver
    is undefined in the reference; here it is refused with a RuntimeError."""
    c = golden["mixed_j6"]
    s2 = pickle.loads(pickle.dumps(_compute(S, c.problem), -1))
    with pytest.raises(RuntimeError):
        s2.predict(c.y, c.xs)


def test_predict_long_series(S):
    from celerite_b200 import synth
    chk = oracle.best()
    N, M = 20000, 3000
    b = synth.make_batch(1, N, 1, 2, seed=8)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 1, 2)
    p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0])
    f = chk.factor(p)
    xs = np.sort(np.random.default_rng(0).uniform(p.x[0] - 1, p.x[-1] + 1, M))
    s = _compute(S, p)
    assert rel_inf(s.predict(b["y"][0], xs), chk.predict(p, f, b["y"][0], xs)) < RTOL_SOLVE


def test_not_computed_raises(S):
    """tests/test_celerite.py:97-100 — RuntimeError before compute."""
    s = S.CholeskySolver()
    assert not s.computed()
    with pytest.raises(RuntimeError):
        s.log_determinant()
    with pytest.raises(RuntimeError):
        s.solve(np.ones(4))
    with pytest.raises(RuntimeError):
        s.dot_solve(np.ones(4))
    with pytest.raises(RuntimeError):
        s.dot_L(np.ones(4))


def test_dimension_mismatch(golden, S):
    c = golden["real_j2"]
    s = _compute(S, c.problem)
    with pytest.raises(RuntimeError):
        s.solve(np.ones(c.problem.N + 1))
    with pytest.raises(RuntimeError):
        s.compute(0.1, [1.0], [1.0, 2.0], [], [], [], [], np.empty(0), np.empty((0, 0)), np.empty((0, 0)),
                  c.problem.x, c.problem.diag)


def test_failed_compute_leaves_solver_uncomputed(golden, S):
    """cholesky.h:57 — computed_ stays false after linalg_exception."""
    s = _compute(S, golden["real_j2"].problem)
    p = golden["npd_j1"].problem
    with pytest.raises(S.LinAlgError):
        s.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp,
                  np.empty(0), np.empty((0, 0)), np.empty((0, 0)), p.x, p.diag)
    assert not s.computed()


def test_pickle_roundtrip(golden, S):
    """tests/test_celerite.py:237-289."""
    s0 = S.CholeskySolver()
    s1 = pickle.loads(pickle.dumps(s0, -1))
    assert not s1.computed()
    c = golden["mixed_j6"]
    s = _compute(S, c.problem)
    s2 = pickle.loads(pickle.dumps(s, -1))
    assert s2.computed() and s2.log_determinant() == s.log_determinant()
    assert rel_inf(s2.solve(c.rhs), c.solve) < RTOL_SOLVE
    assert rel_sum(s2.dot_solve(c.y), c.dot_solve, c.problem.N) < RTOL_SCALAR


def test_recompute_reuses_handle(golden, S):
    s = S.CholeskySolver()
    for name in ("wide_j26", "real_j2", "cfg1_j3", "jitter_only_j0", "complex_j4"):
        c = golden[name]
        p = c.problem
        s.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp,
                  np.empty(0), np.empty((0, 0)), np.empty((0, 0)), p.x, p.diag)
        assert rel_sum(s.log_determinant(), c.log_det, p.N) < RTOL_SCALAR
        assert rel_inf(s.solve(c.rhs), c.solve) < RTOL_SOLVE


@pytest.mark.parametrize("N,Jr,Jc", [(4096, 1, 1), (1000, 0, 2), (20000, 0, 4), (3000, 3, 6), (700, 0, 16)])
def test_random_vs_oracle(S, N, Jr, Jc):
    """Fresh seeded inputs at config-like sizes against the strongest CPU checker available."""
    from celerite_b200 import synth
    chk = oracle.best()
    b = synth.make_batch(1, N, Jr, Jc, seed=100 + N)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], Jr, Jc)
    p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0])
    y = b["y"][0]
    f = chk.factor(p)
    assert f.status == 0
    s = _compute(S, p)
    assert rel_sum(s.log_determinant(), f.log_det, N) < RTOL_SCALAR
    assert rel_sum(s.dot_solve(y), chk.dot_solve(f, y), N) < RTOL_SCALAR
    rhs = np.random.default_rng(N).standard_normal((N, 2))
    assert rel_inf(s.solve(rhs), chk.solve(f, rhs)) < RTOL_SOLVE
    assert rel_inf(s.dot_L(rhs), chk.dot_L(f, rhs)) < RTOL_SOLVE
    out = s.dot(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp,
                np.empty(0), np.empty((0, 0)), np.empty((0, 0)), p.x, rhs)
    assert rel_inf(out, chk.dot(p, rhs)) < RTOL_SOLVE
    # K (K^-1 b) = b with K's diagonal including the measurement variance: dot() leaves it out
    # (cholesky.h:483-485), so add it back
    z = s.solve(rhs)
    kz = s.dot(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp,
               np.empty(0), np.empty((0, 0)), np.empty((0, 0)), p.x, z) + p.diag[:, None] * z
    assert rel_inf(kz, rhs) < 1e-8
    if hasattr(chk, "release"):
        chk.release(f)


def test_factor_file_round_trip_and_residual(tmp_path):
    """audit.save_factor / load_factor carry the pickled state through a file; audit.residual
    checks K (K^-1 y) = y with the stateless dot."""
    from celerite_b200 import audit, solver, synth
    N, Jr, Jc = 3000, 1, 2
    b = synth.make_batch(1, N, Jr, Jc, seed=8)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], Jr, Jc)
    t, d, y = b["t"][0], b["diag"][0], b["y"][0]
    s = solver.CholeskySolver()
    s.compute(j, ar, cr, ac, bc, cc, dc, np.empty(0), np.empty((0, 0)), np.empty((0, 0)), t, d)
    path = str(tmp_path / "factor.npz")
    audit.save_factor(s, path)
    s2 = solver.CholeskySolver()
    state = audit.load_factor(path, s2)
    assert state[1] == N and state[2] == Jr + 2 * Jc and s2.computed()
    assert s2.log_determinant() == s.log_determinant()
    assert np.array_equal(s2.solve(y), s.solve(y))
    assert audit.residual(s, j, ar, cr, ac, bc, cc, dc, t, d, y) < 1e-9


def test_factor_file_crosses_over_to_the_reference_solver_and_back(tmp_path):
    """SURVEY.md §8f-4, both directions: a factor computed on the GPU, written with audit.save_factor, is installed
    in the REFERENCE's CPU CholeskySolver (its own __setstate__, solver.cpp:644-663) and solves there; a factor
    the reference computed and pickled is installed in the GPU solver and solves here.  Same answers to 1e-10."""
    import os
    import pickle
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref_dir = os.path.join(root, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "celerite")):
        pytest.skip("baseline/_ref (pip install of the reference) is not present")
    sys.path.insert(0, ref_dir)
    try:
        from celerite import solver as ref_solver
    finally:
        sys.path.remove(ref_dir)
    from celerite_b200 import audit, solver, synth
    from tests.util import RTOL_SOLVE
    N, Jr, Jc = 2500, 2, 1
    b = synth.make_batch(1, N, Jr, Jc, seed=21)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], Jr, Jc)
    t, d, y = b["t"][0], b["diag"][0], b["y"][0]
    E0, E2 = np.empty(0), np.empty((0, 0))
    gpu = solver.CholeskySolver()
    gpu.compute(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, t, d)
    cpu = ref_solver.CholeskySolver()
    cpu.compute(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, t, d)
    z_cpu, z_gpu = cpu.solve(y), gpu.solve(y)
    # GPU factor -> file -> reference solver
    path = str(tmp_path / "gpu_factor.npz")
    audit.save_factor(gpu, path)
    cpu2 = ref_solver.CholeskySolver.__new__(ref_solver.CholeskySolver)   # what unpickling does (pybind11 py::pickle)
    cpu2.__setstate__(audit.load_factor(path))
    assert cpu2.computed() and abs(cpu2.log_determinant() - cpu.log_determinant()) < 1e-10 * abs(cpu.log_determinant())
    assert rel_inf(np.asarray(cpu2.solve(y)), np.asarray(z_cpu)) < RTOL_SOLVE
    assert abs(cpu2.dot_solve(y) - cpu.dot_solve(y)) < 1e-10 * abs(cpu.dot_solve(y))
    # reference factor -> pickle -> GPU solver
    gpu2 = solver.CholeskySolver()
    gpu2.__setstate__(pickle.loads(pickle.dumps(cpu, -1)).__getstate__())
    assert gpu2.computed() and abs(gpu2.log_determinant() - cpu.log_determinant()) < 1e-10 * abs(cpu.log_determinant())
    assert rel_inf(np.asarray(gpu2.solve(y)), np.asarray(z_cpu)) < RTOL_SOLVE
    assert rel_inf(np.asarray(gpu2.dot_L(y)), np.asarray(cpu.dot_L(y))) < RTOL_SOLVE
    assert rel_inf(np.asarray(z_gpu), np.asarray(z_cpu)) < RTOL_SOLVE
