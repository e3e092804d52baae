"""Shared helpers for the parity tests (test infrastructure)."""
import os

import numpy as np

from oracle import Problem

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_v1.npz")

# Tolerances north_star states (BASELINE.json): 1e-10 relative on log-likelihood,
# log-determinant and solves; 1e-8 relative on gradients; all fp64.
RTOL_SCALAR = 1e-10
RTOL_SOLVE = 1e-10
RTOL_GRAD = 1e-8


def rel_scalar(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def rel_sum(a, b, N, verbose=False):
    """Parity metric for extensive scalars (log det, quadratic form, log-likelihood): the PURE relative
    error |a-b|/|ref| wherever |ref| >= 1 (north_star: "1e-10 relative").  Only a reference value
    below 1 in magnitude -- a sum of N order-one terms that cancelled almost completely, e.g. a
    log-determinant of -0.05 made of 800 logs -- is measured against max(|ref|, N) instead.  With
    verbose=True both figures are printed (pytest -s)."""
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    err = np.abs(a - b)
    pure = err / np.maximum(np.abs(b), 1e-300)
    soft = err / np.maximum(np.abs(b), float(N))
    out = np.where(np.abs(b) >= 1.0, pure, soft)
    if verbose:
        print("rel_sum: pure %.3e  against max(|ref|,N) %.3e  (min |ref| %.3g)"
              % (float(np.max(pure)), float(np.max(soft)), float(np.min(np.abs(b)))))
    return float(np.max(out))


def scalar_parity(got, ref, N, alt=None):
    """(worst error / its tolerance, report string) for extensive scalars.  Error as in rel_sum (pure
    relative for |ref| >= 1).  Tolerance: RTOL_SCALAR, except where the reference disagrees with ITSELF
    by more: `alt` is the same quantity from the other CPU build of the same algorithm (the reference
    compiled with FMA contraction against the plain-C restatement without), and a case whose two CPU
    answers already differ by e (an ill-conditioned kernel) is held to max(1e-10, 3 e)."""
    got, ref = np.atleast_1d(np.asarray(got, dtype=float)), np.atleast_1d(np.asarray(ref, dtype=float))
    den = np.where(np.abs(ref) >= 1.0, np.abs(ref), np.maximum(np.abs(ref), float(N)))
    err = np.abs(got - ref) / den
    tol = np.full(ref.shape, RTOL_SCALAR)
    noise = np.zeros(ref.shape)
    if alt is not None:
        noise = np.abs(np.atleast_1d(np.asarray(alt, dtype=float)) - ref) / den
        tol = np.maximum(tol, 3.0 * noise)
    i = int(np.argmax(err / tol))
    return float(err[i] / tol[i]), "err %.3e tol %.3e (cpu-vs-cpu %.3e, ref %.6g)" % (err[i], tol[i], noise[i], ref[i])


def rel_inf(a, b):
    """||a-b||_inf / ||b||_inf  (SURVEY.md §8c parity metric for vectors)."""
    a, b = np.asarray(a), np.asarray(b)
    den = np.max(np.abs(b)) if b.size else 1.0
    return (np.max(np.abs(a - b)) if b.size else 0.0) / max(den, 1e-300)


def rel_grad(g, ref):
    """per-component |d| / max(|ref|, 1e-12*||ref||_inf), maximised."""
    g, ref = np.asarray(g), np.asarray(ref)
    floor = 1e-12 * max(np.max(np.abs(ref)), 1e-300)
    return np.max(np.abs(g - ref) / np.maximum(np.abs(ref), floor))


class GoldenCase:
    def __init__(self, name, z):
        k = name + "/"
        self.name = name
        g = lambda key: z[k + key] if (k + key) in z.files else None  # noqa: E731
        self.problem = Problem(float(z[k + "jitter"]), g("a_real"), g("c_real"), g("a_comp"), g("b_comp"),
                               g("c_comp"), g("d_comp"), g("x"), g("diag"), A=g("A"), U=g("U"), V=g("V"))
        self.y = g("y")
        self.status = int(z[k + "status"])
        for key in ("phi", "u", "W", "D", "rhs", "solve", "solve_y", "dot_L", "dot", "xs", "predict", "grad"):
            setattr(self, key, g(key))
        for key in ("log_det", "dot_solve", "grad_val"):
            v = g(key)
            setattr(self, key, None if v is None else float(v))
        v = g("grad_status")
        self.grad_status = None if v is None else int(v)


def load_golden():
    z = np.load(GOLDEN)
    return {str(n): GoldenCase(str(n), z) for n in z["names"]}


def golden_names():
    return [str(n) for n in np.load(GOLDEN)["names"]]
