"""Long single series: apply_inverse (solve) and dot_L as chunked scans, on one GPU or with the
time axis split over the GPUs of a node (BASELINE.json config 4).

Single GPU:  ``LongSeries().compute(...)`` then ``.solve(b)`` / ``.dot_L(z)`` with torch CUDA
tensors (N,) or (N, nrhs) — everything stays on the device.

Multi GPU (one process per GPU, ``torch.distributed`` initialised by the caller):
``DistributedLongSeries`` gives rank r the span [lo_r, hi_r) of the series.  Each rank factorises
its own span preceded by a warm-up halo (the factorisation recursion forgets its past like
prod(phi^2), so a halo long enough reproduces the serial factor to round-off), and a sweep is

    summarize (local)  ->  all_gather of (J^2+J) doubles per rank  ->  fold  ->  apply (local)

which is the only communication on the data path.
"""
import ctypes as C

import numpy as np

from . import _capi

OP_SOLVE_FORWARD, OP_DOT_L, OP_SOLVE_BACKWARD = 0, 1, 2


def span_bounds(N, world, rank):
    """Contiguous split of [0, N) into `world` spans; returns (lo, hi) of `rank`."""
    base, rem = divmod(N, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def safe_halo(c_min, dt_mean, tol=1e-18, cap=None):
    """Warm-up length after which a cold-started factorisation agrees with the serial one to
    `tol`: the open-loop bound ln(1/tol) / (2 c_min dt) of SURVEY.md §7.3-3."""
    h = int(np.ceil(np.log(1.0 / tol) / (2.0 * max(c_min * dt_mean, 1e-300))))
    return h if cap is None else min(h, cap)


def fold_carries(summaries, rank, op, nrhs, J):
    """Incoming carry of `rank` from the gathered span summaries.

    summaries: tensor (world, nrhs + nb, J) as written by cel_span_summarize — rows [0, nrhs) are
    the offsets c, rows [nrhs, nrhs+nb) the columns of M (nb = J) or its diagonal (nb = 1, dot_L).
    Forward ops fold ranks 0..rank-1 in order, the backward op folds world-1..rank+1.
    Returns a (nrhs, J) tensor (zeros for the first rank of the walk)."""
    import torch
    world = summaries.shape[0]
    g = torch.zeros((nrhs, J), dtype=summaries.dtype, device=summaries.device)
    order = range(0, rank) if op != OP_SOLVE_BACKWARD else range(world - 1, rank, -1)
    for q in order:
        c = summaries[q, :nrhs, :]
        if op == OP_DOT_L:
            g = g * summaries[q, nrhs, :][None, :] + c
        else:
            M = summaries[q, nrhs:nrhs + J, :].transpose(0, 1)   # M[k][i] = column i, row k
            g = g @ M.transpose(0, 1) + c
    return g.contiguous()


class LongSeries(object):
    """One long series on one GPU (the whole factor is resident)."""

    def __init__(self, device=-1):
        self._lib = _capi.lib()
        self._h = _capi.create(device)
        self.N = 0
        self.J = 0

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None:
            try:
                self._lib.cel_destroy(h)
            except Exception:
                pass

    def compute(self, jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, x, diag, halo=0):
        """cholesky.h:41-210 on host arrays; `halo` leading steps only warm the recursion up."""
        v = lambda a: np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1))  # noqa: E731
        a_real, c_real, a_comp, b_comp, c_comp, d_comp, x, diag = map(v, (a_real, c_real, a_comp, b_comp, c_comp,
                                                                          d_comp, x, diag))
        p = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
        rc = self._lib.cel_compute(self._h, float(jitter), a_real.size, p(a_real), p(c_real), a_comp.size,
                                   p(a_comp), p(b_comp), p(c_comp), p(d_comp), 0, None, None, None,
                                   x.size, p(x), p(diag))
        _capi.check(rc, self._h)
        if halo:
            _capi.check(self._lib.cel_span_trim(self._h, int(halo)), self._h)
        self.N = x.size - int(halo)
        self.J = a_real.size + 2 * a_comp.size

    def _use_stream(self, t):
        import torch
        self._lib.cel_set_stream(self._h, C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream or 1))

    def _cols(self, b):
        import torch
        if not (b.is_cuda and b.dtype == torch.float64):
            raise RuntimeError("expected a float64 CUDA tensor")
        squeeze = b.dim() == 1
        bm = b[:, None] if squeeze else b
        if bm.shape[0] != self.N:
            raise RuntimeError("dimension mismatch")
        # column-major N x nrhs, as the C ABI (and Eigen) lays matrices out
        return bm.transpose(0, 1).contiguous(), squeeze

    def solve(self, b):
        """K^-1 b (GP.apply_inverse, celerite.py:307-328) for a device tensor (N,) or (N, nrhs)."""
        import torch
        cols, squeeze = self._cols(b)
        out = torch.empty_like(cols)
        self._use_stream(b)
        rc = self._lib.cel_solve_device(self._h, cols.shape[0], C.c_void_p(cols.data_ptr()), C.c_void_p(out.data_ptr()))
        _capi.check(rc, self._h)
        out = out.transpose(0, 1)
        return out[:, 0] if squeeze else out

    def dot_L(self, z):
        """L z with K = L L^T (cholesky.h:409-431)."""
        import torch
        cols, squeeze = self._cols(z)
        out = torch.empty_like(cols)
        self._use_stream(z)
        rc = self._lib.cel_dot_L_device(self._h, cols.shape[0], C.c_void_p(cols.data_ptr()), C.c_void_p(out.data_ptr()))
        _capi.check(rc, self._h)
        out = out.transpose(0, 1)
        return out[:, 0] if squeeze else out

    # ---- span primitives (used by DistributedLongSeries) -----------------------------------
    def summarize(self, op, cols):
        import torch
        nrhs = cols.shape[0]
        nb = 1 if op == OP_DOT_L else self.J
        out = torch.empty((nrhs + nb, self.J), dtype=torch.float64, device=cols.device)
        self._use_stream(cols)
        rc = self._lib.cel_span_summarize(self._h, op, nrhs, C.c_void_p(cols.data_ptr()), C.c_void_p(out.data_ptr()))
        _capi.check(rc, self._h)
        return out

    def apply(self, op, cols, carry_in, reuse_summary=True):
        import torch
        out = torch.empty_like(cols)
        self._use_stream(cols)
        cptr = C.c_void_p(carry_in.data_ptr()) if carry_in is not None else None
        rc = self._lib.cel_span_apply(self._h, op, cols.shape[0], C.c_void_p(cols.data_ptr()), cptr,
                                      C.c_void_p(out.data_ptr()), 1 if reuse_summary else 0)
        _capi.check(rc, self._h)
        return out

    def log_determinant(self, device=None):
        """sum of log D over the span's own steps, as a 0-d device tensor (cholesky.h:208)."""
        import torch
        out = torch.empty((), dtype=torch.float64, device=device if device is not None else torch.device("cuda", torch.cuda.current_device()))
        self._use_stream(out)
        _capi.check(self._lib.cel_span_log_determinant(self._h, C.c_void_p(out.data_ptr())), self._h)
        return out

    def quad(self, xd):
        """sum (x/D)^2 D over the span for xd = the result of a forward-solve sweep (1, N) or (N,)."""
        import torch
        xd = xd.reshape(-1).contiguous()
        if xd.shape[0] != self.N:
            raise RuntimeError("dimension mismatch")
        out = torch.empty((), dtype=torch.float64, device=xd.device)
        self._use_stream(xd)
        _capi.check(self._lib.cel_span_quad(self._h, C.c_void_p(xd.data_ptr()), C.c_void_p(out.data_ptr())), self._h)
        return out

    def dot_solve(self, y):
        """y^T K^-1 y (cholesky.h:326-401) for a device vector: forward sweep + ordered reduction."""
        cols, _ = self._cols(y)
        return self.quad(self.apply(OP_SOLVE_FORWARD, cols, None, reuse_summary=False))

    def log_likelihood(self, y):
        """-0.5 (y^T K^-1 y + log det K + N log 2 pi) as a 0-d device tensor (celerite.py:214-216)."""
        import math
        return -0.5 * (self.dot_solve(y) + self.log_determinant(y.device) + self.N * math.log(2.0 * math.pi))

    def fold(self, op, gathered, nrhs, world, rank):
        """Incoming carry of `rank` from the gathered summaries (world, nrhs + nb, J), on the device
        (one small kernel; `fold_carries` is the same arithmetic in torch)."""
        import torch
        out = torch.empty((nrhs, self.J), dtype=torch.float64, device=gathered.device)
        self._use_stream(gathered)
        rc = self._lib.cel_span_fold(self._h, op, nrhs, world, rank, C.c_void_p(gathered.data_ptr()),
                                     C.c_void_p(out.data_ptr()))
        _capi.check(rc, self._h)
        return out

    def launch_count(self):
        return int(self._lib.cel_launch_count(self._h))


class DistributedLongSeries(object):
    """One long series split by time over the ranks of ``torch.distributed`` (one GPU each).

    On CUDA the whole exchange lives in the C library (cel_comm_init / cel_*_mgpu: summarize ->
    ncclAllGather -> fold -> apply enqueued from C++ on one stream, one host call per operation);
    torch.distributed only carries the 128-byte NCCL id from rank 0 to the others once.  With
    ``use_c_abi=False`` (and in the CPU tests of the host logic, gloo backend) the all-gather goes
    through torch.distributed instead and the fold is the torch statement `fold_carries`."""

    def __init__(self, device=-1, group=None, use_c_abi=True):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.local = LongSeries(device)
        self.lo = self.hi = self.halo = 0
        self.c_abi = False
        self.collective_backend = "torch.distributed all_gather_into_tensor"
        if use_c_abi and self.world > 1:
            self._init_comm()

    def _init_comm(self):
        import torch
        lib, h = self.local._lib, self.local._h
        dev = torch.device("cuda", torch.cuda.current_device())
        backend = self.dist.get_backend(self.group)
        ident = torch.zeros(_capi.COMM_ID_BYTES, dtype=torch.uint8)
        if self.rank == 0:
            buf = (C.c_ubyte * _capi.COMM_ID_BYTES)()
            _capi.check(lib.cel_comm_unique_id(buf), h)
            ident = torch.frombuffer(bytearray(buf), dtype=torch.uint8).clone()
        if backend == "nccl":
            ident = ident.to(dev)
        src = self.dist.get_global_rank(self.group, 0) if self.group is not None else 0
        self.dist.broadcast(ident, src=src, group=self.group)
        raw = bytes(ident.cpu().numpy().tobytes())
        _capi.check(lib.cel_comm_init(h, self.world, self.rank, C.c_char_p(raw)), h)
        self.c_abi = True
        self.collective_backend = "ncclAllGather / ncclAllReduce inside the C ABI (cel_*_mgpu)"

    def compute(self, jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, x, diag, halo):
        """x, diag: the FULL host arrays (every rank slices its own span + halo)."""
        N = len(x)
        self.lo, self.hi = span_bounds(N, self.world, self.rank)
        self.halo = min(int(halo), self.lo)
        if self.c_abi:
            v = lambda a: np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1))  # noqa: E731
            a_real, c_real, a_comp, b_comp, c_comp, d_comp, x, diag = map(v, (a_real, c_real, a_comp, b_comp, c_comp,
                                                                              d_comp, x, diag))
            p = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
            lib, h = self.local._lib, self.local._h
            rc = lib.cel_compute_mgpu(h, float(jitter), a_real.size, p(a_real), p(c_real), a_comp.size, p(a_comp),
                                      p(b_comp), p(c_comp), p(d_comp), N, p(x), p(diag), int(halo))
            _capi.check(rc, h)
            self.local.N = self.hi - self.lo
            self.local.J = a_real.size + 2 * a_comp.size
            self.N_total = N
            return
        s0 = self.lo - self.halo
        self.local.compute(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, x[s0:self.hi], diag[s0:self.hi],
                           halo=self.halo)
        self.N_total = N

    def _sweep(self, op, cols):
        import torch
        summ = self.local.summarize(op, cols)
        # concatenated along dim 0 (the form both NCCL and gloo accept), viewed per rank afterwards
        gathered = torch.empty((self.world * summ.shape[0], summ.shape[1]), dtype=summ.dtype, device=summ.device)
        self.dist.all_gather_into_tensor(gathered, summ, group=self.group)
        if gathered.is_cuda:
            carry = self.local.fold(op, gathered, cols.shape[0], self.world, self.rank)
        else:
            gathered = gathered.view(self.world, summ.shape[0], summ.shape[1])
            carry = fold_carries(gathered, self.rank, op, cols.shape[0], self.local.J)
        return self.local.apply(op, cols, carry, reuse_summary=True)

    def _mgpu(self, fn, b_local):
        import torch
        cols, squeeze = self.local._cols(b_local)
        out = torch.empty_like(cols)
        self.local._use_stream(b_local)
        h = self.local._h
        _capi.check(fn(h, cols.shape[0], C.c_void_p(cols.data_ptr()), C.c_void_p(out.data_ptr())), h)
        out = out.transpose(0, 1)
        return out[:, 0] if squeeze else out

    def solve(self, b_local):
        if self.c_abi:
            return self._mgpu(self.local._lib.cel_solve_mgpu, b_local)
        cols, squeeze = self.local._cols(b_local)
        xd = self._sweep(OP_SOLVE_FORWARD, cols)
        z = self._sweep(OP_SOLVE_BACKWARD, xd)
        z = z.transpose(0, 1)
        return z[:, 0] if squeeze else z

    def dot_L(self, z_local):
        if self.c_abi:
            return self._mgpu(self.local._lib.cel_dot_L_mgpu, z_local)
        cols, squeeze = self.local._cols(z_local)
        out = self._sweep(OP_DOT_L, cols).transpose(0, 1)
        return out[:, 0] if squeeze else out

    def log_determinant(self, device=None):
        """log det K of the whole series: every rank's partial, added with one all-reduce."""
        ld = self.local.log_determinant(device).reshape(1)
        self.dist.all_reduce(ld, group=self.group)
        return ld[0]

    def dot_solve(self, y_local):
        """y^T K^-1 y of the whole series: one forward sweep, partial sums added with one all-reduce."""
        cols, _ = self.local._cols(y_local)
        q = self.local.quad(self._sweep(OP_SOLVE_FORWARD, cols)).reshape(1)
        self.dist.all_reduce(q, group=self.group)
        return q[0]

    def log_likelihood(self, y_local, N_total=None):
        """celerite.py:214-216 for the whole series (N_total = its length over all ranks)."""
        import math
        if self.c_abi:
            cols, _ = self.local._cols(y_local)
            self.local._use_stream(y_local)
            out = (C.c_double * 3)()
            h = self.local._h
            _capi.check(self.local._lib.cel_log_likelihood_mgpu(h, C.c_void_p(cols.data_ptr()), out), h)
            return float(out[2])
        N_total = self.N_total if N_total is None else N_total
        return -0.5 * (self.dist_solve_logdet(y_local) + N_total * math.log(2.0 * math.pi))

    def dist_solve_logdet(self, y_local):
        return self.dot_solve(y_local) + self.log_determinant(y_local.device)


class CapturedSweep(object):
    """A sweep of a LongSeries (one GPU) recorded once as a CUDA graph and replayed.

    ``solve`` / ``dot_L`` on a long series are 6-12 short launches with host calls in between.
    Everything in them is stream-ordered and allocation-free once the handle's buffers have grown,
    so the whole call can be captured (useful for short series, where launch latency shows; at
    N = 1e7 it measured 2.54 ms against 2.51 ms plain).  Not for DistributedLongSeries: capturing
    the all-gather hung NCCL on the 2-GPU box it was tried on.

        f = CapturedSweep(eng, "solve", example_rhs)     # records after three warm-up calls
        z = f(rhs)                                        # copies rhs into the static input, replays

    The result is a static tensor that the next call overwrites.  Re-record after ``compute``.
    """

    def __init__(self, engine, method, example, warmup=3):
        import torch
        if isinstance(engine, DistributedLongSeries):
            raise RuntimeError("CapturedSweep records single-GPU sweeps only")
        self.engine = engine
        fn = getattr(engine, method)
        self.inp = example.clone()
        local = getattr(engine, "local", engine)
        side = torch.cuda.Stream(device=example.device)
        side.wait_stream(torch.cuda.current_stream(example.device))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                fn(self.inp)
        torch.cuda.current_stream(example.device).wait_stream(side)
        torch.cuda.synchronize(example.device)
        n0 = local.launch_count()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = fn(self.inp)
        self.launches = local.launch_count() - n0      # kernels of this library inside one replay

    def __call__(self, x):
        self.inp.copy_(x)
        self.graph.replay()
        return self.out
