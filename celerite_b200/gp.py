"""Batched GP front door on kernel PARAMETERS.

The reference evaluates one series per call: ``GP.log_likelihood`` (celerite.py:180-219) and
``GP.grad_log_likelihood`` (celerite.py:221-297), the kernel's parameters being turned into
coefficient arrays by ``Term.get_all_coefficients`` (terms.py:156-188) and the coefficient
gradient being chained to the parameters with an autograd Jacobian (terms.py:197-215).  Here a
whole stack of parameter vectors (MCMC walkers, or one kernel per light curve) is mapped to
coefficient rows on the device (csrc/termmaps.cu), swept in one launch, and the gradient comes
back with respect to the reference's own parameters.

    spec = TermSum(["sho", "real", "jitter"])          # same order as term1 + term2 + term3
    gp = BatchedGP(spec)
    ll = gp.log_likelihood(params, t, y, yerr)           # params: (B, spec.n_params) torch CUDA
    ll, grad = gp.grad_log_likelihood(params, t, y, yerr)

Everything stays on the device: inputs and outputs are float64 torch CUDA tensors.
"""
import ctypes as C

import numpy as np

from . import _capi
from .batched import BatchedSolver

KINDS = {"jitter": 0, "real": 1, "complex3": 2, "complex4": 3, "sho": 4, "sho_overdamped": 5, "matern32": 6,
         "product": 7, "sum": 8}
_NAMES = {
    0: ("log_sigma",), 1: ("log_a", "log_c"), 2: ("log_a", "log_c", "log_d"),
    3: ("log_a", "log_b", "log_c", "log_d"), 4: ("log_S0", "log_Q", "log_omega0"),
    5: ("log_S0", "log_Q", "log_omega0"), 6: ("log_sigma", "log_rho"),
}


class TermSum(object):
    """A sum of celerite terms, by kind (terms.py:280-324).  An entry ``("product", x, y)`` is the product
    ``x * y`` (terms.py:218-277) of two operands, each a kind, another product, or ``("sum", x, y)`` for
    ``(x + y)``; every coefficient set of ``x`` meets every set of ``y`` as in the reference.  The kind list
    may also be given already flattened to prefix order (7 = product, 8 = sum, each followed by its two
    operands).  ``eps`` maps a position in the flattened list to the Matern32Term approximation parameter
    (default 0.01, terms.py:549)."""

    def __init__(self, kinds, eps=None):
        def flat(k):
            if isinstance(k, (tuple, list)):
                if len(k) != 3 or k[0] not in ("product", "sum"):
                    raise ValueError('a product is written ("product", x, y), a sum inside it ("sum", x, y)')
                return [KINDS[k[0]]] + flat(k[1]) + flat(k[2])
            if isinstance(k, str):
                if k not in KINDS:
                    raise ValueError("unknown term kind %r" % (k,))
                return [KINDS[k]]
            return [int(k)]
        out = []
        for k in kinds:
            out += flat(k)
        self.kinds = np.array(out, dtype=np.int32)
        self.eps = np.full(len(self.kinds), 0.01)
        for i, v in (eps or {}).items():
            self.eps[i] = v
        npar, jr, jc = C.c_int(), C.c_int(), C.c_int()
        rc = _capi.lib().cel_term_layout(len(self.kinds), C.c_void_p(self.kinds.ctypes.data), C.byref(npar),
                                         C.byref(jr), C.byref(jc))
        if rc != _capi.OK:
            raise ValueError("unknown term kind, malformed product / sum, jitter inside a product, or too many entries")
        self.n_params, self.Jr, self.Jc = npar.value, jr.value, jc.value
        self.n_coeffs = 1 + 2 * self.Jr + 4 * self.Jc

    @property
    def parameter_names(self):
        """Names in the reference's convention (modeling.py): ``terms[i]:log_x`` for the i-th term of the
        sum, ``terms[i]:k1:log_x`` / ``:k2:`` inside a product, ``:terms[j]:`` inside a sum."""
        names = []
        kinds = [int(k) for k in self.kinds]

        def walk(pos, prefix):
            k = kinds[pos]
            if k == 7:
                pos = walk(pos + 1, prefix + "k1:")
                return walk(pos, prefix + "k2:")
            if k == 8:
                pos = walk(pos + 1, prefix + "terms[0]:")
                return walk(pos, prefix + "terms[1]:")
            names.extend(prefix + n for n in _NAMES[k])
            return pos + 1

        pos = node = 0
        while pos < len(kinds):
            pos = walk(pos, "terms[%d]:" % node)
            node += 1
        return tuple(names)

    def _args(self):
        return len(self.kinds), C.c_void_p(self.kinds.ctypes.data), C.c_void_p(self.eps.ctypes.data)


class BatchedGP(object):
    """log-likelihood (and gradient) of B series under B parameter vectors of one term sum."""

    def __init__(self, spec, device=-1):
        self.spec = spec
        self.solver = BatchedSolver(device)
        self._lib = self.solver._lib
        self._h = self.solver._h

    def _use_stream(self, x):
        import torch
        self._lib.cel_set_stream(self._h, C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream or 1))

    def _check(self, params):
        import torch
        if not (params.is_cuda and params.dtype == torch.float64 and params.is_contiguous() and params.dim() == 2):
            raise RuntimeError("params: contiguous float64 CUDA tensor of shape (B, n_params)")
        if params.shape[1] != self.spec.n_params:
            raise RuntimeError("dimension mismatch: %d parameters per series expected" % self.spec.n_params)

    def coefficients(self, params):
        """(B, n_params) -> coefficient rows (B, 1+2Jr+4Jc) and a per-series status (1: an SHO
        term's Q is on the other side of 1/2 than the spec says)."""
        import torch
        self._check(params)
        B = params.shape[0]
        rows = torch.empty((B, self.spec.n_coeffs), dtype=torch.float64, device=params.device)
        status = torch.empty(B, dtype=torch.int32, device=params.device)
        self._use_stream(params)
        n, kinds, eps = self.spec._args()
        rc = self._lib.cel_term_coefficients(self._h, n, kinds, eps, B, C.c_void_p(params.data_ptr()), params.shape[1],
                                             C.c_void_p(rows.data_ptr()), rows.shape[1], C.c_void_p(status.data_ptr()))
        _capi.check(rc, self._h)
        return rows, status

    def backprop(self, params, grad_coeffs):
        """Coefficient gradient (B, 1+2Jr+4Jc) -> parameter gradient (B, n_params)."""
        import torch
        self._check(params)
        B = params.shape[0]
        out = torch.empty((B, self.spec.n_params), dtype=torch.float64, device=params.device)
        self._use_stream(params)
        n, kinds, eps = self.spec._args()
        rc = self._lib.cel_term_backprop(self._h, n, kinds, eps, B, C.c_void_p(params.data_ptr()), params.shape[1],
                                         C.c_void_p(grad_coeffs.data_ptr()), grad_coeffs.shape[1],
                                         C.c_void_p(out.data_ptr()), out.shape[1])
        _capi.check(rc, self._h)
        return out

    @staticmethod
    def _diag(yerr, t):
        import torch
        if yerr is None:
            return torch.zeros_like(t)
        return (yerr * yerr).contiguous()                     # celerite.py:141: diag = yerr**2

    def log_likelihood(self, params, t, y, yerr=None, lengths=None):
        """-0.5*(r^T K^-1 r + log det K + N log 2 pi) per series (celerite.py:214-216); -inf where
        the matrix is not positive definite (``quiet=True``, celerite.py:203-208).  ``lengths``
        (int32 CUDA tensor, optional): series b uses only its first lengths[b] points."""
        rows, _ = self.coefficients(params)
        return self.solver.log_likelihood(rows, t, self._diag(yerr, t), y, self.spec.Jr, self.spec.Jc,
                                          lengths=lengths)

    def grad_log_likelihood(self, params, t, y, yerr=None):
        """(ll, grad) per series, grad with respect to the parameters (celerite.py:221-297).
        ll is the value ``log_likelihood`` returns; the reference's own value at this call adds
        pi*ln(N) instead of N*ln(2 pi) (solver.cpp:434), which the drop-in
        single-series entry point reproduces and this batched one does not."""
        import math
        rows, _ = self.coefficients(params)
        val, g, status = self.solver.grad_log_likelihood(rows, t, self._diag(yerr, t), y, self.spec.Jr, self.spec.Jc)
        N = t.shape[-1]
        return val - 0.5 * N * math.log(2.0 * math.pi), self.backprop(params, g)
