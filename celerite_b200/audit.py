"""Factor serialisation and a residual checker.

* ``save_factor`` / ``load_factor``: the solver state the reference pickles (solver.cpp:36-58:
  computed, N, J, log_det, phi, u, W, D) as one ``.npz`` file with a format tag, so a factor
  computed on the GPU can be audited elsewhere (or by the reference's CPU solver, whose
  ``__setstate__`` takes the same 8-tuple).
* ``residual``: ||K z - y||_inf / ||y||_inf through the stateless ``dot`` (cholesky.h:444-590)
  plus the measurement variances — the check of paper/figures/error/error.py:46-59, which does
  not depend on the factor it audits.
"""
import numpy as np

FORMAT = "celerite_b200.factor.v1"
E0, E2 = np.empty(0), np.empty((0, 0))


def save_factor(solver, path):
    comp, n, j, ld, phi, u, W, D = solver.__getstate__()
    np.savez_compressed(path, format=np.array(FORMAT), computed=np.array(bool(comp)), N=np.array(n), J=np.array(j),
                        log_det=np.array(ld), phi=phi, u=u, W=W, D=D)


def load_factor(path, solver=None):
    """Returns the 8-tuple state; with ``solver`` given, also installs it (``__setstate__``)."""
    z = np.load(path)
    if str(z["format"]) != FORMAT:
        raise ValueError("not a %s file" % FORMAT)
    state = (bool(z["computed"]), int(z["N"]), int(z["J"]), float(z["log_det"]), np.asfortranarray(z["phi"]),
             np.asfortranarray(z["u"]), np.asfortranarray(z["W"]), np.ascontiguousarray(z["D"]))
    if solver is not None:
        solver.__setstate__(state)
    return state


def residual(solver, jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, x, diag, y, z=None):
    """Relative residual of z ~ K^-1 y (z = solver.solve(y) when not given)."""
    y = np.asarray(y, dtype=np.float64)
    if z is None:
        z = solver.solve(y)
    z = np.asarray(z, dtype=np.float64).reshape(len(y), -1)[:, 0]
    kz = solver.dot(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, E0, E2, E2, x, z)
    kz = np.asarray(kz).reshape(len(y), -1)[:, 0] + np.asarray(diag) * z
    return float(np.max(np.abs(kz - y)) / np.max(np.abs(y)))
