"""Batched front door: B independent series (and kernels) per call.

No counterpart in the reference, whose Python API is one series per call
(celerite/celerite.py:180-219); this is what an MCMC / survey user of that path needs on a GPU.
Inputs may be

* numpy arrays (host): the C library stages them through the device in chunks, overlapping the
  copies with the sweeps, and returns numpy arrays; or
* torch CUDA tensors (device-resident): pointers are passed straight through, outputs are torch
  tensors on the same device, and the work is queued on torch's current stream.

Layout: ``t, diag, y`` are (B, N) row-major (or 1-D (N,) to share one series across the batch);
``coeffs`` is (B, P) or (P,) with rows ``[jitter, a_real[Jr], c_real[Jr], a_comp[Jc], b_comp[Jc],
c_comp[Jc], d_comp[Jc]]`` — the order of the reference's gradient vector (solver.cpp:392-406).
"""
import ctypes as C

import numpy as np

from . import _capi


def _is_torch(x):
    return type(x).__module__.startswith("torch")


class BatchedSolver(object):
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

    def launch_count(self):
        return int(self._lib.cel_launch_count(self._h))

    # ------------------------------------------------------------------------------------------
    def _shapes(self, coeffs, t, Jr, Jc):
        P = 1 + 2 * Jr + 4 * Jc
        if t.ndim == 1:
            N, Bt, stride = t.shape[0], None, 0
        else:
            Bt, N = t.shape
            stride = N
        if coeffs.ndim == 1:
            Bc, cstride = None, 0
        else:
            Bc, cstride = coeffs.shape[0], coeffs.shape[1]
        if coeffs.shape[-1] < P:
            raise RuntimeError("dimension mismatch: coefficient rows need 1+2*Jr+4*Jc entries")
        if Bt is not None and Bc is not None and Bt != Bc:
            raise RuntimeError("dimension mismatch: batch sizes differ")
        B = Bt if Bt is not None else (Bc if Bc is not None else 1)
        return B, N, stride, cstride

    def log_likelihood(self, coeffs, t, diag, y, Jr, Jc, return_parts=False, lengths=None):
        """-0.5*(y^T K^-1 y + log det K + N log 2pi) per series (celerite.py:214-216).

        ``lengths`` (B int32, optional) makes the batch ragged: series b has ``lengths[b]`` points
        and the rest of its row is padding (any finite values); N in the constant is then the
        series' own length.

        Returns ``ll`` (and ``logdet, quad, status`` when ``return_parts``).  Series whose
        factorisation met D_n < 0 (cholesky.h:176) get ll = -inf and status = 3, the batched
        equivalent of ``GP.log_likelihood(quiet=True)`` (celerite.py:203-218).
        """
        if _is_torch(t):
            return self._ll_torch(coeffs, t, diag, y, Jr, Jc, return_parts, lengths)
        coeffs = np.ascontiguousarray(coeffs, dtype=np.float64)
        t = np.ascontiguousarray(t, dtype=np.float64)
        diag = np.ascontiguousarray(diag, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.float64)
        if diag.shape != t.shape or y.shape != t.shape:
            raise RuntimeError("dimension mismatch")
        B, N, stride, cstride = self._shapes(coeffs, t, Jr, Jc)
        ll = np.empty(B); logdet = np.empty(B); quad = np.empty(B)
        status = np.empty(B, dtype=np.int32)
        p = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
        if lengths is None:
            rc = self._lib.cel_batched_log_likelihood(self._h, B, N, stride, Jr, Jc, p(coeffs), cstride,
                                                      p(t), p(diag), p(y), p(logdet), p(quad), p(ll), p(status),
                                                      _capi.HOST_POINTERS)
        else:
            lengths = np.ascontiguousarray(lengths, dtype=np.int32)
            if lengths.shape != (B,):
                raise RuntimeError("dimension mismatch: one length per series")
            rc = self._lib.cel_batched_log_likelihood_ragged(self._h, B, N, stride, Jr, Jc, p(coeffs), cstride,
                                                             p(t), p(diag), p(y), p(lengths), p(logdet), p(quad),
                                                             p(ll), p(status), _capi.HOST_POINTERS)
        _capi.check(rc, self._h)
        ll[~np.isfinite(ll)] = -np.inf
        return (ll, logdet, quad, status) if return_parts else ll

    def _ll_torch(self, coeffs, t, diag, y, Jr, Jc, return_parts, lengths=None):
        import torch
        for a in (coeffs, t, diag, y):
            if not (a.is_cuda and a.dtype == torch.float64 and a.is_contiguous()):
                raise RuntimeError("expected contiguous float64 CUDA tensors")
        if diag.shape != t.shape or y.shape != t.shape:
            raise RuntimeError("dimension mismatch")
        B, N, stride, cstride = self._shapes(coeffs, t, Jr, Jc)
        out = torch.empty((3, B), dtype=torch.float64, device=t.device)
        status = torch.empty(B, dtype=torch.int32, device=t.device)
        # torch's default stream has handle 0; the C ABI reads NULL as "use your own stream", so
        # name the legacy default stream explicitly (cudaStreamLegacy == 0x1)
        self._lib.cel_set_stream(self._h, C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream or 1))
        p = lambda a: C.c_void_p(a.data_ptr())  # noqa: E731
        if lengths is None:
            rc = self._lib.cel_batched_log_likelihood(self._h, B, N, stride, Jr, Jc, p(coeffs), cstride,
                                                      p(t), p(diag), p(y), p(out[0]), p(out[1]), p(out[2]),
                                                      p(status), 0)
        else:
            if not (lengths.is_cuda and lengths.dtype == torch.int32 and lengths.is_contiguous() and lengths.shape == (B,)):
                raise RuntimeError("lengths: contiguous int32 CUDA tensor of shape (B,)")
            rc = self._lib.cel_batched_log_likelihood_ragged(self._h, B, N, stride, Jr, Jc, p(coeffs), cstride,
                                                             p(t), p(diag), p(y), p(lengths), p(out[0]), p(out[1]),
                                                             p(out[2]), p(status), 0)
        _capi.check(rc, self._h)
        ll = torch.where(torch.isfinite(out[2]), out[2], torch.full_like(out[2], float("-inf")))
        return (ll, out[0], out[1], status) if return_parts else ll

    def grad_log_likelihood_ragged(self, coeffs, t, diag, y, lengths, Jr, Jc):
        """As ``grad_log_likelihood`` for a ragged batch: torch CUDA tensors only, ``t, diag, y`` of
        shape (B, N) with series b occupying the first ``lengths[b]`` entries of its row."""
        import torch
        for a in (coeffs, t, diag, y):
            if not (a.is_cuda and a.dtype == torch.float64 and a.is_contiguous()):
                raise RuntimeError("expected contiguous float64 CUDA tensors")
        if t.dim() != 2 or diag.shape != t.shape or y.shape != t.shape:
            raise RuntimeError("dimension mismatch")
        B, N, stride, cstride = self._shapes(coeffs, t, Jr, Jc)
        if not (lengths.is_cuda and lengths.dtype == torch.int32 and lengths.is_contiguous() and lengths.shape == (B,)):
            raise RuntimeError("lengths: contiguous int32 CUDA tensor of shape (B,)")
        Pn = 1 + 2 * Jr + 4 * Jc
        val = torch.empty(B, dtype=torch.float64, device=t.device)
        grad = torch.empty((B, Pn), dtype=torch.float64, device=t.device)
        status = torch.empty(B, dtype=torch.int32, device=t.device)
        self._lib.cel_set_stream(self._h, C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream or 1))
        p = lambda a: C.c_void_p(a.data_ptr())  # noqa: E731
        rc = self._lib.cel_batched_grad_log_likelihood_ragged(self._h, B, N, stride, Jr, Jc, p(coeffs), cstride, p(t),
                                                              p(diag), p(y), p(lengths), p(val), p(grad), p(status))
        _capi.check(rc, self._h)
        return val, grad, status

    def grad_log_likelihood(self, coeffs, t, diag, y, Jr, Jc):
        """Per series: val = -0.5*(y^T K^-1 y + log det K) (no constant) and its gradient with
        respect to the coefficient row, in the layout of the reference's gradient vector
        (solver.cpp:392-434, incl. the zeroed jitter slot for jitter <= DBL_EPSILON).
        Returns (val[B], grad[B, P], status[B]); numpy in -> numpy out, torch CUDA in -> torch out."""
        Pn = 1 + 2 * Jr + 4 * Jc
        if _is_torch(t):
            import torch
            for a in (coeffs, t, diag, y):
                if not (a.is_cuda and a.dtype == torch.float64 and a.is_contiguous()):
                    raise RuntimeError("expected contiguous float64 CUDA tensors")
            B, N, stride, cstride = self._shapes(coeffs, t, Jr, Jc)
            val = torch.empty(B, dtype=torch.float64, device=t.device)
            grad = torch.empty((B, Pn), dtype=torch.float64, device=t.device)
            status = torch.empty(B, dtype=torch.int32, device=t.device)
            self._lib.cel_set_stream(self._h, C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream or 1))
            p = lambda a: C.c_void_p(a.data_ptr())  # noqa: E731
            flags = 0
        else:
            coeffs = np.ascontiguousarray(coeffs, dtype=np.float64)
            t = np.ascontiguousarray(t, dtype=np.float64)
            diag = np.ascontiguousarray(diag, dtype=np.float64)
            y = np.ascontiguousarray(y, dtype=np.float64)
            B, N, stride, cstride = self._shapes(coeffs, t, Jr, Jc)
            val = np.empty(B); grad = np.empty((B, Pn)); status = np.empty(B, dtype=np.int32)
            p = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
            flags = _capi.HOST_POINTERS
        if diag.shape != t.shape or y.shape != t.shape:
            raise RuntimeError("dimension mismatch")
        rc = self._lib.cel_batched_grad_log_likelihood(self._h, B, N, stride, Jr, Jc, p(coeffs), cstride,
                                                       p(t), p(diag), p(y), p(val), p(grad), p(status), flags)
        _capi.check(rc, self._h)
        return val, grad, status

    def compute(self, coeffs, t, diag, Jr, Jc, lengths=None):
        """Batched CholeskySolver::compute (cholesky.h:41-210) on torch CUDA tensors: returns
        (phi, u, W, D, logdet, status) with phi,u: (B, N-1, J), W: (B, N, J), D: (B, N) — per
        series the reference's factor layout (J fastest).  ``lengths`` (int32, shape (B,)): a ragged
        batch, series b being the first lengths[b] entries of its row; its factor is valid in those
        columns, the log-determinant and the status count its own steps only."""
        import torch
        for a in (coeffs, t, diag):
            if not (a.is_cuda and a.dtype == torch.float64 and a.is_contiguous()):
                raise RuntimeError("expected contiguous float64 CUDA tensors")
        B, N, stride, cstride = self._shapes(coeffs, t, Jr, Jc)
        if lengths is not None and not (lengths.is_cuda and lengths.dtype == torch.int32 and lengths.is_contiguous()
                                        and lengths.shape == (B,)):
            raise RuntimeError("lengths: contiguous int32 CUDA tensor of shape (B,)")
        J = Jr + 2 * Jc
        dev = t.device
        phi = torch.empty((B, max(N - 1, 0), J), dtype=torch.float64, device=dev)
        u = torch.empty((B, max(N - 1, 0), J), dtype=torch.float64, device=dev)
        W = torch.empty((B, N, J), dtype=torch.float64, device=dev)
        D = torch.empty((B, N), dtype=torch.float64, device=dev)
        logdet = torch.empty(B, dtype=torch.float64, device=dev)
        status = torch.empty(B, dtype=torch.int32, device=dev)
        self._lib.cel_set_stream(self._h, C.c_void_p(torch.cuda.current_stream(dev).cuda_stream or 1))
        p = lambda a: C.c_void_p(a.data_ptr())  # noqa: E731
        if lengths is not None:
            rc = self._lib.cel_batched_compute_ragged(self._h, B, N, stride, Jr, Jc, p(coeffs), cstride, p(t), p(diag),
                                                      p(lengths), p(phi), p(u), p(W), p(D), p(logdet), p(status))
        else:
            rc = self._lib.cel_batched_compute(self._h, B, N, stride, Jr, Jc, p(coeffs), cstride, p(t), p(diag),
                                               p(phi), p(u), p(W), p(D), p(logdet), p(status))
        _capi.check(rc, self._h)
        return phi, u, W, D, logdet, status

    def raw_log_likelihood(self, B, N, stride, Jr, Jc, coeffs_ptr, cstride, t_ptr, diag_ptr, y_ptr,
                           logdet_ptr, quad_ptr, ll_ptr, status_ptr, flags=0, stream=None):
        """Thin pass-through of cel_batched_log_likelihood for callers that manage memory
        themselves (bench.py): integers are raw device (or host, with HOST_POINTERS) addresses."""
        if stream is not None:
            self._lib.cel_set_stream(self._h, C.c_void_p(stream or 1))  # 0 -> cudaStreamLegacy
        v = lambda x: C.c_void_p(x) if x else None  # noqa: E731
        rc = self._lib.cel_batched_log_likelihood(self._h, B, N, stride, Jr, Jc, v(coeffs_ptr), cstride,
                                                  v(t_ptr), v(diag_ptr), v(y_ptr), v(logdet_ptr), v(quad_ptr),
                                                  v(ll_ptr), v(status_ptr), flags)
        _capi.check(rc, self._h)
