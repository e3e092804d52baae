"""TEST INFRASTRUCTURE — ctypes front-ends for the CPU oracle.

Two back-ends with one interface:

* ``Port``      — oracle/libcelerite_oracle.so, the plain-C restatement (celerite_oracle.c)
* ``Reference`` — oracle/_ref/libceleref.so, the UNMODIFIED reference headers behind a thin
                  extern "C" harness (ref_harness.cpp); built only where /root/reference is
                  mounted, shipped prebuilt to the GPU box.

Only tests/, ``__graft_entry__.smoke()`` and bench.py's CPU-baseline legs may import this
package.  The product path (celerite_b200) must never do so.

Arrays: ``phi,u`` have shape (N-1, J) C-order == J x (N-1) column-major (J fastest, as in
cholesky.h:76-78); ``W`` is (N, J); right-hand sides are (N, nrhs) Fortran order.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def build(verbose=False):
    """Run oracle/Makefile (C port always; the reference harness when /root/reference exists)."""
    r = subprocess.run(["make", "-C", _HERE, "all"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed")


def _p(a):
    if a is None:
        return None
    return a.ctypes.data_as(_dp)


def _v(a):
    return np.ascontiguousarray(np.atleast_1d(np.asarray(a, dtype=np.float64)))


def _f(a, N):
    a = np.asarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a[:, None]
    assert a.shape[0] == N
    return np.asfortranarray(a)


class Problem:
    """One series + one kernel, in the reference's argument order (cholesky.h:41-54)."""

    def __init__(self, jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, x, diag,
                 A=None, U=None, V=None):
        self.jitter = float(jitter)
        self.a_real, self.c_real = _v(a_real), _v(c_real)
        self.a_comp, self.b_comp = _v(a_comp), _v(b_comp)
        self.c_comp, self.d_comp = _v(c_comp), _v(d_comp)
        self.x, self.diag = _v(x), _v(diag)
        self.N = self.x.shape[0]
        self.Jr, self.Jc = self.a_real.shape[0], self.a_comp.shape[0]
        if U is not None and np.size(U) > 0:
            self.U = np.asfortranarray(np.asarray(U, dtype=np.float64))  # (Jg, N), Jg fastest
            self.V = np.asfortranarray(np.asarray(V, dtype=np.float64))
            self.A = _v(A)
            self.Jg = self.U.shape[0]
        else:
            self.U = self.V = self.A = None
            self.Jg = 0
        self.J = self.Jr + 2 * self.Jc + self.Jg
        self.P = 1 + 2 * self.Jr + 4 * self.Jc

    def coeff_args(self):
        return (C.c_double(self.jitter),
                C.c_int(self.Jr), _p(self.a_real), _p(self.c_real),
                C.c_int(self.Jc), _p(self.a_comp), _p(self.b_comp), _p(self.c_comp), _p(self.d_comp),
                C.c_int(self.Jg), _p(self.A), _p(self.U), _p(self.V))

    def coeff_row(self):
        """[jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp] (solver.cpp:392-406)."""
        return np.concatenate([[self.jitter], self.a_real, self.c_real, self.a_comp,
                               self.b_comp, self.c_comp, self.d_comp])


class Factor:
    def __init__(self, J, N):
        self.J, self.N = J, N
        self.phi = np.zeros((max(N - 1, 0), J))
        self.u = np.zeros((max(N - 1, 0), J))
        self.W = np.zeros((N, J))
        self.D = np.zeros(N)
        self.log_det = float("nan")
        self.status = -1


class Port:
    kind = "port"

    def __init__(self):
        path = os.path.join(_HERE, "libcelerite_oracle.so")
        if not os.path.exists(path):
            build()
        self.lib = C.CDLL(path)
        self.lib.celo_dot_solve.restype = C.c_double

    def factor(self, p):
        f = Factor(p.J, p.N)
        ld = C.c_double(float("nan"))
        f.status = self.lib.celo_factor(*p.coeff_args(), C.c_int(p.N), _p(p.x), _p(p.diag),
                                        _p(f.phi), _p(f.u), _p(f.W), _p(f.D), C.byref(ld))
        f.log_det = ld.value
        return f

    def dot_solve(self, f, b):
        b = _v(b)
        return self.lib.celo_dot_solve(C.c_int(f.J), C.c_int(f.N), _p(f.phi), _p(f.u), _p(f.W), _p(f.D), _p(b))

    def solve(self, f, b):
        b = _f(b, f.N)
        out = np.zeros_like(b, order="F")
        self.lib.celo_solve(C.c_int(f.J), C.c_int(f.N), _p(f.phi), _p(f.u), _p(f.W), _p(f.D),
                            C.c_int(b.shape[1]), _p(b), _p(out))
        return out

    def dot_L(self, f, z):
        z = _f(z, f.N)
        out = np.zeros_like(z, order="F")
        self.lib.celo_dot_L(C.c_int(f.J), C.c_int(f.N), _p(f.phi), _p(f.u), _p(f.W), _p(f.D),
                            C.c_int(z.shape[1]), _p(z), _p(out))
        return out

    def dot(self, p, z):
        z = _f(z, p.N)
        out = np.zeros_like(z, order="F")
        self.lib.celo_dot(*p.coeff_args(), C.c_int(p.N), _p(p.x), C.c_int(z.shape[1]), _p(z), _p(out))
        return out

    def predict(self, p, f, y, xs):
        y, xs = _v(y), _v(xs)
        out = np.zeros(xs.shape[0])
        self.lib.celo_predict(C.c_int(p.Jr), _p(p.a_real), _p(p.c_real),
                              C.c_int(p.Jc), _p(p.a_comp), _p(p.b_comp), _p(p.c_comp), _p(p.d_comp),
                              C.c_int(f.J), C.c_int(f.N), _p(p.x), _p(f.phi), _p(f.u), _p(f.W), _p(f.D),
                              _p(y), C.c_int(xs.shape[0]), _p(xs), _p(out))
        return out

    def grad(self, p, y):
        y = _v(y)
        val = C.c_double(float("nan"))
        g = np.zeros(p.P)
        st = self.lib.celo_grad(*p.coeff_args(), C.c_int(p.N), _p(p.x), _p(y), _p(p.diag), C.byref(val), _p(g))
        return st, val.value, g

    def batched_loglik(self, B, N, stride, Jr, Jc, coeffs, coeff_stride, t, diag, y, nthreads=1):
        out = np.zeros((B, 3))
        self.lib.celo_batched_loglik(C.c_long(B), C.c_int(N), C.c_long(stride), C.c_int(Jr), C.c_int(Jc),
                                     _p(coeffs), C.c_long(coeff_stride), _p(t), _p(diag), _p(y), _p(out))
        return out


class Reference:
    kind = "reference"

    @staticmethod
    def path():
        return os.path.join(_HERE, "_ref", "libceleref.so")

    @classmethod
    def available(cls):
        return os.path.exists(cls.path())

    def __init__(self):
        if not self.available():
            build()
        self.lib = C.CDLL(self.path())
        self.lib.celref_create.restype = C.c_void_p

    def hardware_threads(self):
        return int(self.lib.celref_hardware_threads())

    def _with_solver(self, p):
        h = C.c_void_p(self.lib.celref_create())
        st = self.lib.celref_compute(h, *p.coeff_args(), C.c_int(p.N), _p(p.x), _p(p.diag))
        return h, st

    def factor(self, p):
        f = Factor(p.J, p.N)
        h, st = self._with_solver(p)
        f.status = st
        f._h = h
        if st == 0:
            ld = C.c_double()
            self.lib.celref_log_determinant(h, C.byref(ld))
            f.log_det = ld.value
            self.lib.celref_get_factor(h, _p(f.phi), _p(f.u), _p(f.W), _p(f.D))
        return f

    def release(self, f):
        if getattr(f, "_h", None) is not None:
            self.lib.celref_destroy(f._h)
            f._h = None

    def dot_solve(self, f, b):
        b = _v(b)
        out = C.c_double()
        st = self.lib.celref_dot_solve(f._h, C.c_int(b.shape[0]), _p(b), C.byref(out))
        assert st == 0, st
        return out.value

    def solve(self, f, b):
        b = _f(b, f.N)
        out = np.zeros_like(b, order="F")
        st = self.lib.celref_solve(f._h, C.c_int(f.N), C.c_int(b.shape[1]), _p(b), _p(out))
        assert st == 0, st
        return out

    def dot_L(self, f, z):
        z = _f(z, f.N)
        out = np.zeros_like(z, order="F")
        st = self.lib.celref_dot_L(f._h, C.c_int(f.N), C.c_int(z.shape[1]), _p(z), _p(out))
        assert st == 0, st
        return out

    def dot(self, p, z):
        z = _f(z, p.N)
        out = np.zeros_like(z, order="F")
        st = self.lib.celref_dot(*p.coeff_args(), C.c_int(p.N), _p(p.x), C.c_int(z.shape[1]), _p(z), _p(out))
        assert st == 0, st
        return out

    def predict(self, p, f, y, xs):
        y, xs = _v(y), _v(xs)
        out = np.zeros(xs.shape[0])
        st = self.lib.celref_predict(f._h, C.c_int(f.N), _p(y), C.c_int(xs.shape[0]), _p(xs), _p(out))
        assert st == 0, st
        return out

    def grad(self, p, y):
        y = _v(y)
        val = C.c_double(float("nan"))
        g = np.zeros(p.P)
        st = self.lib.celref_grad(*p.coeff_args(), C.c_int(p.N), _p(p.x), _p(y), _p(p.diag), C.byref(val), _p(g))
        return st, val.value, g

    def batched_loglik(self, B, N, stride, Jr, Jc, coeffs, coeff_stride, t, diag, y, nthreads=1):
        out = np.zeros((B, 3))
        self.lib.celref_batched_loglik(C.c_long(B), C.c_int(N), C.c_long(stride), C.c_int(Jr), C.c_int(Jc),
                                       _p(coeffs), C.c_long(coeff_stride), _p(t), _p(diag), _p(y), _p(out),
                                       C.c_int(nthreads))
        return out

    def batched_grad(self, B, N, stride, Jr, Jc, coeffs, coeff_stride, t, diag, y, nthreads=1):
        P = 1 + 2 * Jr + 4 * Jc
        val = np.zeros(B)
        g = np.zeros((B, P))
        st = np.zeros(B, dtype=np.int32)
        self.lib.celref_batched_grad(C.c_long(B), C.c_int(N), C.c_long(stride), C.c_int(Jr), C.c_int(Jc),
                                     _p(coeffs), C.c_long(coeff_stride), _p(t), _p(diag), _p(y),
                                     _p(val), _p(g), st.ctypes.data_as(_ip), C.c_int(nthreads))
        return val, g, st


def best():
    """The strongest checker available: the compiled reference if present, else the C port."""
    return Reference() if Reference.available() else Port()
