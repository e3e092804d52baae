"""Python mirror of the reference's ``celerite.solver`` extension module (celerite/solver.cpp),
backed by the CUDA library through the C ABI.

Same class name, method names, argument order and return shapes as the pybind11 module the
reference builds, so that ``celerite.GP`` (celerite/celerite.py) runs on it unchanged:

    solver.CholeskySolver().compute(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp,
                                    A, U, V, x, diag)                 solver.cpp:467-483
    .solve(b) -> (N, nrhs) Fortran-ordered, always 2-D                 solver.cpp:507-509
    .dot_solve(b) -> float                                             solver.cpp:527-529
    .dot_L(z), .dot(...12 args...), .predict(y, x)                     solver.cpp:541-590
    .log_determinant(), .computed()                                    solver.cpp:594-611
    .grad_log_likelihood(...13 args...) -> (float, ndarray)            solver.cpp:347-463
    .__getstate__/__setstate__ 8-tuple                                 solver.cpp:36-58, 644-663
    LinAlgError, has_autodiff(), get_library_version()                 solver.cpp:76-87
"""
import ctypes as C
import math

import numpy as np

from . import _capi
from ._capi import LinAlgError  # noqa: F401  (re-exported like solver.LinAlgError)

__all__ = ["CholeskySolver", "LinAlgError", "has_autodiff", "get_library_version"]

_DBL_EPSILON = 2.220446049250313e-16


def get_library_version():
    return _capi.lib().cel_version().decode()


def has_autodiff():
    """The reference answers True when built with AutoDiff; here gradients come from a
    hand-written reverse-mode sweep and are always available."""
    return True


def _vec(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1))


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _mat(b, N=None):
    """numpy -> N x nrhs column-major, as pybind's Eigen caster does (1-D becomes N x 1)."""
    b = np.asarray(b, dtype=np.float64)
    if b.ndim == 1:
        b = b[:, None]
    if b.ndim != 2:
        raise RuntimeError("expected a vector or a matrix")
    return np.asfortranarray(b)


class _Coeffs:
    def __init__(self, jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, N):
        self.jitter = float(jitter)
        self.a_real, self.c_real = _vec(a_real), _vec(c_real)
        self.a_comp, self.b_comp = _vec(a_comp), _vec(b_comp)
        self.c_comp, self.d_comp = _vec(c_comp), _vec(d_comp)
        # cholesky.h:60-63
        if self.a_real.size != self.c_real.size or not (
                self.a_comp.size == self.b_comp.size == self.c_comp.size == self.d_comp.size):
            raise RuntimeError("dimension mismatch")
        A = _vec(A) if A is not None else np.empty(0)
        U = np.asarray(U, dtype=np.float64) if U is not None else np.empty((0, 0))
        V = np.asarray(V, dtype=np.float64) if V is not None else np.empty((0, 0))
        if U.ndim != 2:
            U = U.reshape(0, 0) if U.size == 0 else np.atleast_2d(U)
        if V.ndim != 2:
            V = V.reshape(0, 0) if V.size == 0 else np.atleast_2d(V)
        has_general = A.size != 0
        # cholesky.h:65-69
        if has_general and (A.size != N or U.shape[1] != N or V.shape[1] != N):
            raise RuntimeError("dimension mismatch")
        if U.shape[0] != V.shape[0]:
            raise RuntimeError("dimension mismatch")
        self.A = A if has_general else None
        self.Jg = int(U.shape[0])
        self.U = np.asfortranarray(U) if self.Jg else None
        self.V = np.asfortranarray(V) if self.Jg else None
        self.Jr, self.Jc = int(self.a_real.size), int(self.a_comp.size)

    def args(self):
        return (self.jitter, self.Jr, _ptr(self.a_real), _ptr(self.c_real),
                self.Jc, _ptr(self.a_comp), _ptr(self.b_comp), _ptr(self.c_comp), _ptr(self.d_comp),
                self.Jg, _ptr(self.A), _ptr(self.U), _ptr(self.V))


class CholeskySolver(object):
    """Drop-in for ``celerite.solver.CholeskySolver`` running on one B200."""

    def __init__(self, device=-1):
        self._lib = _capi.lib()
        self._h = _capi.create(device)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None:
            try:
                self._lib.cel_destroy(h)
            except Exception:
                pass

    # -- compute ------------------------------------------------------------------------------
    def compute(self, jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, x, diag):
        x, diag = _vec(x), _vec(diag)
        N = x.size
        if diag.size != N:
            raise RuntimeError("dimension mismatch")  # cholesky.h:59
        c = _Coeffs(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, N)
        rc = self._lib.cel_compute(self._h, *c.args(), N, _ptr(x), _ptr(diag))
        _capi.check(rc, self._h)

    def computed(self):
        return bool(self._lib.cel_computed(self._h))

    def log_determinant(self):
        out = C.c_double()
        _capi.check(self._lib.cel_log_determinant(self._h, C.byref(out)), self._h)
        return out.value

    # -- sweeps -------------------------------------------------------------------------------
    def dot_solve(self, b):
        b = _vec(b)
        out = C.c_double()
        _capi.check(self._lib.cel_dot_solve(self._h, b.size, _ptr(b), C.byref(out)), self._h)
        return out.value

    def solve(self, b):
        b = _mat(b)
        out = np.empty_like(b, order="F")
        _capi.check(self._lib.cel_solve(self._h, b.shape[0], b.shape[1], _ptr(b), _ptr(out)), self._h)
        return out

    def dot_L(self, z):
        z = _mat(z)
        out = np.empty_like(z, order="F")
        _capi.check(self._lib.cel_dot_L(self._h, z.shape[0], z.shape[1], _ptr(z), _ptr(out)), self._h)
        return out

    def dot(self, jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, x, b):
        x = _vec(x)
        b = _mat(b)
        if b.shape[0] != x.size:
            raise RuntimeError("dimension mismatch")  # cholesky.h:459
        c = _Coeffs(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, x.size)
        out = np.empty_like(b, order="F")
        rc = self._lib.cel_dot(self._h, *c.args(), x.size, _ptr(x), b.shape[1], _ptr(b), _ptr(out))
        _capi.check(rc, self._h)
        return out

    def predict(self, y, x):
        y, x = _vec(y), _vec(x)
        out = np.empty(x.size)
        _capi.check(self._lib.cel_predict(self._h, y.size, _ptr(y), x.size, _ptr(x), _ptr(out)), self._h)
        return out

    # -- gradient -----------------------------------------------------------------------------
    def grad_log_likelihood(self, jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, x, y, diag):
        x, y, diag = _vec(x), _vec(y), _vec(diag)
        N = x.size
        if diag.size != N or y.size != N:
            raise RuntimeError("dimension mismatch")
        c = _Coeffs(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, N)
        grad = np.zeros(1 + 2 * c.Jr + 4 * c.Jc)
        val = C.c_double()
        rc = self._lib.cel_grad_log_likelihood(self._h, *c.args(), N, _ptr(x), _ptr(y), _ptr(diag),
                                               C.byref(val), _ptr(grad))
        _capi.check(rc, self._h)
        # solver.cpp:415 — the reference's constant is M_PI*log(N), not N*log(2*pi)
        return val.value - 0.5 * math.pi * math.log(N), grad

    # -- pickle (solver.cpp:36-58) ------------------------------------------------------------
    def __getstate__(self):
        comp, N, J = C.c_int(), C.c_int(), C.c_int()
        ld = C.c_double()
        _capi.check(self._lib.cel_get_state(self._h, C.byref(comp), C.byref(N), C.byref(J), C.byref(ld),
                                            None, None, None, None), self._h)
        n, j = N.value, J.value
        phi = np.zeros((j, max(n - 1, 0)), order="F")
        u = np.zeros((j, max(n - 1, 0)), order="F")
        W = np.zeros((j, n), order="F")
        D = np.zeros(n)
        if comp.value:
            _capi.check(self._lib.cel_get_state(self._h, None, None, None, None,
                                                _ptr(phi), _ptr(u), _ptr(W), _ptr(D)), self._h)
        return (bool(comp.value), n, j, ld.value, phi, u, W, D)

    def __setstate__(self, state):
        if len(state) != 8:
            raise RuntimeError("Invalid state!")  # solver.cpp:645
        if getattr(self, "_h", None) is None:
            self._lib = _capi.lib()
            self._h = _capi.create(-1)
        comp, n, j, ld, phi, u, W, D = state
        phi = np.asfortranarray(phi, dtype=np.float64)
        u = np.asfortranarray(u, dtype=np.float64)
        W = np.asfortranarray(W, dtype=np.float64)
        D = _vec(D)
        n, j = int(n), int(j)
        # cel_set_state copies j*(n-1), j*(n-1), j*n and n doubles from these pointers: a truncated or
        # corrupted state must end here, not in an out-of-bounds read
        if n < 0 or j < 0 or phi.shape != (j, max(n - 1, 0)) or u.shape != phi.shape \
                or W.shape != (j, n) or D.size != n:
            raise RuntimeError("Invalid state!")
        rc = self._lib.cel_set_state(self._h, int(bool(comp)), int(n), int(j), float(ld),
                                     _ptr(phi), _ptr(u), _ptr(W), _ptr(D))
        _capi.check(rc, self._h)

    # -- extras -------------------------------------------------------------------------------
    def launch_count(self):
        return int(self._lib.cel_launch_count(self._h))
