"""GPU parity, batched fused log-likelihood (K1) vs the CPU oracle on identical seeded inputs."""
import numpy as np
import pytest

import oracle
from tests.util import RTOL_SCALAR, rel_inf, rel_sum, scalar_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def bs():
    from celerite_b200.batched import BatchedSolver
    return BatchedSolver()


def _check(bs, B, N, Jr, Jc, seed, torch_path=False):
    from celerite_b200 import synth
    b = synth.make_batch(B, N, Jr, Jc, seed=seed)
    P = b["coeffs"].shape[1]
    ref = oracle.best().batched_loglik(B, N, N, Jr, Jc, b["coeffs"], P, b["t"], b["diag"], b["y"], nthreads=8)
    assert np.all(ref[:, 2] == 0)
    if torch_path:
        import torch
        dev = torch.device("cuda:0")
        tt = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
        ll, logdet, quad, status = bs.log_likelihood(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"]),
                                                     Jr, Jc, return_parts=True)
        torch.cuda.synchronize()
        ll, logdet, quad, status = (x.cpu().numpy() for x in (ll, logdet, quad, status))
    else:
        ll, logdet, quad, status = bs.log_likelihood(b["coeffs"], b["t"], b["diag"], b["y"], Jr, Jc,
                                                     return_parts=True)
    assert np.all(status == 0)
    # the other CPU build of the same algorithm gives the reference's own round-off floor per series
    alt = oracle.Port().batched_loglik(B, N, N, Jr, Jc, b["coeffs"], P, b["t"], b["diag"], b["y"], nthreads=8)
    c = 0.5 * N * np.log(2 * np.pi)
    for got, r, a in ((logdet, ref[:, 0], alt[:, 0]), (quad, ref[:, 1], alt[:, 1]),
                      (ll, -0.5 * (ref[:, 1] + ref[:, 0]) - c, -0.5 * (alt[:, 1] + alt[:, 0]) - c)):
        worst, msg = scalar_parity(got, r, N, a)
        assert worst <= 1.0, msg


@pytest.mark.parametrize("B,N,Jr,Jc", [
    (300, 2000, 0, 1),    # cfg2 unit (J=2), B not a multiple of the CTA size
    (130, 1000, 0, 2),    # cfg5 unit (J=4)
    (64, 4096, 1, 1),     # cfg1 kernel (J=3)
    (33, 777, 2, 0),      # real only, odd N (unaligned rows: the non-TMA path)
    (17, 501, 1, 2),      # J=5
    (40, 1500, 0, 3),     # J=6
    (9, 1203, 0, 4),      # J=8  (cooperative kernel, 8 lanes per series)
    (5, 900, 2, 5),       # J=12
    (3, 800, 0, 8),       # J=16
    (3, 640, 4, 11),      # J=26 (the widest kernel of tests/test_celerite.py)
    (2, 700, 0, 16),      # J=32
    (3, 300, 4, 18),      # J=40: wider than one row per lane, the wide kernels (widej.cu)
    (2, 200, 0, 40),      # J=80
])
def test_batched_loglik_host(bs, B, N, Jr, Jc):
    _check(bs, B, N, Jr, Jc, seed=B + N)


@pytest.mark.parametrize("B,N,Jr,Jc", [(257, 2000, 0, 1), (10, 333, 1, 1), (4, 1000, 0, 8)])
def test_batched_loglik_device(bs, B, N, Jr, Jc):
    _check(bs, B, N, Jr, Jc, seed=7 * B + N, torch_path=True)


def test_edge_sizes(bs):
    for (B, N) in [(1, 1), (1, 2), (2, 7), (65, 8), (3, 9), (1, 16), (1, 17)]:
        _check(bs, B, N, 0, 1, seed=B * 100 + N)


def test_shared_data_many_kernels(bs):
    """MCMC-walker layout: one light curve, many kernels (stride 0 on the data)."""
    from celerite_b200 import synth
    B, N, Jr, Jc = 50, 800, 0, 2
    b = synth.make_batch(B, N, Jr, Jc, seed=5)
    t, d, y = b["t"][0], b["diag"][0], b["y"][0]
    ll = bs.log_likelihood(b["coeffs"], t, d, y, Jr, Jc)
    T, Dg, Y = (np.repeat(a[None], B, 0) for a in (t, d, y))
    ref = oracle.best().batched_loglik(B, N, N, Jr, Jc, b["coeffs"], b["coeffs"].shape[1], T, Dg, Y)
    ll_ref = -0.5 * (ref[:, 1] + ref[:, 0] + N * np.log(2 * np.pi))
    assert rel_inf(ll, ll_ref) < RTOL_SCALAR


def test_shared_kernel_many_series(bs):
    from celerite_b200 import synth
    B, N, Jr, Jc = 40, 600, 1, 1
    b = synth.make_batch(B, N, Jr, Jc, seed=6)
    row = b["coeffs"][0].copy()
    ll = bs.log_likelihood(row, b["t"], b["diag"], b["y"], Jr, Jc)
    C = np.repeat(row[None], B, 0)
    ref = oracle.best().batched_loglik(B, N, N, Jr, Jc, C, C.shape[1], b["t"], b["diag"], b["y"])
    ll_ref = -0.5 * (ref[:, 1] + ref[:, 0] + N * np.log(2 * np.pi))
    assert rel_inf(ll, ll_ref) < RTOL_SCALAR


def test_not_positive_definite_series_is_flagged_not_fatal(bs):
    """cholesky.h:176 per series: the bad series reports status 3 / -inf, the others are fine."""
    from celerite_b200 import synth
    B, N = 70, 300
    b = synth.make_batch(B, N, 1, 0, seed=11)
    b["coeffs"][13, 0] = 0.0
    b["coeffs"][13, 1] = -1.0          # negative amplitude
    b["diag"][13] = 1e-3
    ll, logdet, quad, status = bs.log_likelihood(b["coeffs"], b["t"], b["diag"], b["y"], 1, 0, return_parts=True)
    ref = oracle.best().batched_loglik(B, N, N, 1, 0, b["coeffs"], 3, b["t"], b["diag"], b["y"])
    assert ref[13, 2] == 3 and status[13] == 3 and ll[13] == -np.inf
    ok = np.arange(B) != 13
    assert np.all(status[ok] == 0)
    assert rel_sum(logdet[ok], ref[ok, 0], N) < RTOL_SCALAR


def test_full_size_properties(bs):
    """cfg2 at full per-series size on a slice of the batch: size-independent properties.
    (1) duplicating the batch gives identical results per copy (determinism / no cross-talk);
    (2) scaling y by 2 scales the quadratic form by 4 and leaves log det unchanged."""
    import torch
    from celerite_b200 import synth
    B, N = 4096, 2000
    b = synth.make_batch(B, N, 0, 1, seed=2)
    dev = torch.device("cuda:0")
    tt = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    c, t, d, y = tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"])
    ll1, ld1, q1, st1 = bs.log_likelihood(c, t, d, y, 0, 1, return_parts=True)
    ll2, ld2, q2, st2 = bs.log_likelihood(c, t, d, 2.0 * y, 0, 1, return_parts=True)
    torch.cuda.synchronize()
    assert int(st1.sum()) == 0
    assert torch.equal(ld1, ld2)
    assert float(((q2 - 4 * q1).abs() / q1.abs()).max()) < 1e-13
    idx = torch.randperm(B, device=dev)
    ll3 = bs.log_likelihood(c[idx].contiguous(), t[idx].contiguous(), d[idx].contiguous(), y[idx].contiguous(), 0, 1)
    torch.cuda.synchronize()
    assert torch.equal(ll3, ll1[idx])
    # and a sample of series against the oracle
    pick = np.arange(0, B, 257)
    ref = oracle.best().batched_loglik(len(pick), N, N, 0, 1, np.ascontiguousarray(b["coeffs"][pick]), 5,
                                       np.ascontiguousarray(b["t"][pick]), np.ascontiguousarray(b["diag"][pick]),
                                       np.ascontiguousarray(b["y"][pick]))
    assert rel_sum(ld1.cpu().numpy()[pick], ref[:, 0], N) < RTOL_SCALAR
    assert rel_sum(q1.cpu().numpy()[pick], ref[:, 1], N) < RTOL_SCALAR


def test_repeatable_at_full_size(bs):
    """Regression test for a cross-proxy hazard in the TMA staging ring (generic-proxy reads of a
    slot vs the async-proxy refill): the cfg2-sized launch must give bit-identical results on
    every repetition."""
    import torch
    from celerite_b200 import synth
    B, N = 65536, 2000
    rng = np.random.Generator(np.random.PCG64(4))
    c = synth.make_coeffs(rng, B, 0, 1)
    t = np.empty((B, N)); d = np.empty((B, N)); y = np.empty((B, N))
    synth.fill_series(rng, t, d, y)
    dev = torch.device("cuda:0")
    tt = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    C, T, D, Y = tt(c), tt(t), tt(d), tt(y)
    ll0, ld0, q0, st0 = bs.log_likelihood(C, T, D, Y, 0, 1, return_parts=True)
    torch.cuda.synchronize()
    assert int(st0.sum()) == 0
    for _ in range(25):
        ll, ld, q, st = bs.log_likelihood(C, T, D, Y, 0, 1, return_parts=True)
        torch.cuda.synchronize()
        assert torch.equal(ld, ld0) and torch.equal(q, q0) and torch.equal(st, st0)


@pytest.mark.parametrize("B,N,Jr,Jc", [(100, 500, 0, 1), (70, 403, 0, 2), (40, 300, 0, 3), (33, 200, 1, 1),
                                       (5, 333, 0, 1), (3, 150, 0, 8), (3, 120, 3, 17)])
def test_batched_compute_materialises_reference_factor(bs, B, N, Jr, Jc):
    """cel_batched_compute: phi,u,W,D of every series vs the oracle's factor (the TMA-store
    lockstep kernel for all-complex even J <= 6 with aligned rows, the cooperative kernel else)."""
    import torch
    from celerite_b200 import synth
    from tests.util import RTOL_SOLVE
    b = synth.make_batch(B, N, Jr, Jc, seed=3 * B + N)
    dev = torch.device("cuda:0")
    tt = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    phi, u, W, D, logdet, status = bs.compute(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), Jr, Jc)
    torch.cuda.synchronize()
    assert int(status.sum()) == 0
    chk = oracle.best()
    for i in list(range(0, B, max(1, B // 7))) + [B - 1]:
        j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][i], Jr, Jc)
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][i], b["diag"][i])
        f = chk.factor(p)
        assert rel_inf(phi[i].cpu().numpy(), f.phi) < RTOL_SOLVE
        assert rel_inf(u[i].cpu().numpy(), f.u) < RTOL_SOLVE
        assert rel_inf(W[i].cpu().numpy(), f.W) < RTOL_SOLVE
        assert rel_inf(D[i].cpu().numpy(), f.D) < RTOL_SOLVE
        assert rel_sum(float(logdet[i]), f.log_det, N) < RTOL_SCALAR
        if hasattr(chk, "release"):
            chk.release(f)


@pytest.mark.parametrize("B,N,Jr,Jc", [(40, 300, 0, 1), (21, 257, 1, 2), (9, 200, 2, 7)])
def test_ragged_batched_compute(bs, B, N, Jr, Jc):
    """cel_batched_compute_ragged: every series is factorised over its own length whatever its padding holds
    (a negative diagonal there included); log-determinant and status count the series' own steps; lengths
    N, 1 and 2 among them.  Factor columns beyond a series' length are unspecified and not looked at."""
    import torch
    from celerite_b200 import synth
    from tests.util import RTOL_SOLVE
    b = synth.make_batch(B, N, Jr, Jc, seed=7 * B + N)
    rng = np.random.default_rng(B + N)
    lengths = rng.integers(3, N + 1, B).astype(np.int32)
    lengths[0], lengths[1], lengths[2] = N, 1, 2
    for i in range(B):
        b["diag"][i, lengths[i]:] = -1.0e3           # padding that would fail the factorisation if it counted
        b["t"][i, lengths[i]:] = 0.0                 # ... and is not even sorted
    dev = torch.device("cuda:0")
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    phi, u, W, D, logdet, status = bs.compute(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), Jr, Jc, lengths=tt(lengths))
    torch.cuda.synchronize()
    assert int(status.sum()) == 0
    chk = oracle.best()
    for i in range(B):
        n = int(lengths[i])
        j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][i], Jr, Jc)
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][i, :n].copy(), b["diag"][i, :n].copy())
        f = chk.factor(p)
        if n > 1:
            assert rel_inf(phi[i, :n - 1].cpu().numpy(), f.phi) < RTOL_SOLVE
            assert rel_inf(u[i, :n - 1].cpu().numpy(), f.u) < RTOL_SOLVE
        assert rel_inf(W[i, :n].cpu().numpy(), f.W) < RTOL_SOLVE
        assert rel_inf(D[i, :n].cpu().numpy(), f.D) < RTOL_SOLVE
        assert rel_sum(float(logdet[i]), f.log_det, n) < RTOL_SCALAR
        if hasattr(chk, "release"):
            chk.release(f)


def test_wide_kernels_batched_gradient_and_limit(bs):
    """J = 37 (3 real + 17 complex): value and all 1 + 2*3 + 4*17 derivatives of a small batch against the
    reference's forward-mode AD; and the width the library refuses (J > 96)."""
    from celerite_b200 import synth, _capi
    from tests.util import RTOL_GRAD, rel_grad
    B, N, Jr, Jc = 3, 150, 3, 17
    b = synth.make_batch(B, N, Jr, Jc, seed=91)
    P = b["coeffs"].shape[1]
    val, grad, status = bs.grad_log_likelihood(b["coeffs"], b["t"], b["diag"], b["y"], Jr, Jc)
    gv, gg, gs = oracle.best().batched_grad(B, N, N, Jr, Jc, b["coeffs"], P, b["t"], b["diag"], b["y"], nthreads=4)
    assert np.all(status == 0) and np.all(gs == 0)
    const = 0.5 * np.pi * np.log(N)                      # solver.cpp:415, carried by the reference's value only
    assert np.max(np.abs((val - const) - gv) / np.abs(gv)) < RTOL_SCALAR
    for i in range(B):
        assert rel_grad(grad[i], gg[i]) < RTOL_GRAD
    wide = synth.make_batch(1, 16, 0, 49, seed=1)        # J = 98
    with pytest.raises(_capi.CeleriteB200Error):
        bs.log_likelihood(wide["coeffs"], wide["t"], wide["diag"], wide["y"], 0, 49)


def test_unusual_series_take_the_careful_path(bs):
    """The second-generation fused kernel is optimistic: pivots outside [2^-500, 2^500] and trigonometric
    arguments beyond 1e9 send a lane (resp. a warp) to careful_series().  Both must still match the reference:
    a batch in which some series are scaled by 1e160 (pivots ~1e160), and one whose time stamps are offset
    by 1e9 days (d*t ~ 1e10).  The same inputs run through the ragged entry point (first-generation kernel)."""
    from celerite_b200 import synth
    B, N, Jr, Jc = 70, 400, 0, 1
    b = synth.make_batch(B, N, Jr, Jc, seed=77)
    coeffs, t, diag, y = b["coeffs"].copy(), b["t"].copy(), b["diag"].copy(), b["y"].copy()
    big = [3, 40, 69]
    for i in big:                       # K -> s K, y -> sqrt(s) y: same quadratic form, log det + N log s
        s = 1e160
        coeffs[i, 0] *= s               # jitter
        coeffs[i, 1] *= s               # a
        coeffs[i, 2] *= s               # b
        diag[i] *= s
        y[i] *= 1e80
    P = coeffs.shape[1]
    ref = oracle.best().batched_loglik(B, N, N, Jr, Jc, coeffs, P, t, diag, y, nthreads=4)
    ll, logdet, quad, status = bs.log_likelihood(coeffs, t, diag, y, Jr, Jc, return_parts=True)
    assert np.all(status == 0) and np.all(ref[:, 2] == 0)
    assert np.max(np.abs(logdet - ref[:, 0]) / np.abs(ref[:, 0])) < RTOL_SCALAR
    assert np.max(np.abs(quad - ref[:, 1]) / np.abs(ref[:, 1])) < RTOL_SCALAR
    assert logdet[3] > 0.9 * N * np.log(1e160)                    # the scaled series really are huge
    # time stamps far from zero: sin/cos of d*t with |d*t| ~ 1e10 (stock range reduction on the GPU as on the CPU)
    t2 = t + 1.0e9
    ref2 = oracle.best().batched_loglik(B, N, N, Jr, Jc, b["coeffs"], P, t2, b["diag"], b["y"], nthreads=4)
    ll2, logdet2, quad2, status2 = bs.log_likelihood(b["coeffs"], t2, b["diag"], b["y"], Jr, Jc, return_parts=True)
    assert np.all(status2 == 0)
    assert np.max(np.abs(logdet2 - ref2[:, 0]) / np.maximum(np.abs(ref2[:, 0]), 1.0)) < 1e-9
    assert np.max(np.abs(quad2 - ref2[:, 1]) / np.abs(ref2[:, 1])) < 1e-9


def test_jitter_only_batch(bs):
    """J = 0 (cholesky.h:90-95, 334-339): K = diag(sigma^2 + jitter)."""
    rng = np.random.default_rng(0)
    B, N = 37, 100
    diag = rng.uniform(0.1, 0.5, (B, N)) ** 2
    y = rng.standard_normal((B, N))
    t = np.sort(rng.uniform(0, 10, (B, N)), axis=1)
    coeffs = rng.uniform(0.01, 0.1, (B, 1))
    ll, logdet, quad, status = bs.log_likelihood(coeffs, t, diag, y, 0, 0, return_parts=True)
    D = diag + coeffs
    assert np.all(status == 0)
    assert rel_inf(logdet, np.log(D).sum(1)) < RTOL_SCALAR
    assert rel_inf(quad, (y * y / D).sum(1)) < RTOL_SCALAR


def test_padded_rows_device_and_host(bs):
    """stride > N (rows padded to a fixed pitch), on both the device and the host-pointer path."""
    import ctypes as C
    import torch
    from celerite_b200 import _capi, synth
    B, N, pitch, Jr, Jc = 70, 333, 400, 0, 1
    b = synth.make_batch(B, N, Jr, Jc, seed=21)
    pad = lambda a: np.ascontiguousarray(np.pad(a, ((0, 0), (0, pitch - N)), constant_values=np.nan))  # noqa: E731
    T, D, Y = pad(b["t"]), pad(b["diag"]), pad(b["y"])
    ref = oracle.best().batched_loglik(B, N, N, Jr, Jc, b["coeffs"], 5, b["t"], b["diag"], b["y"])
    lib = bs._lib
    p = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
    ld = np.empty(B); q = np.empty(B); ll = np.empty(B); st = np.empty(B, dtype=np.int32)
    rc = lib.cel_batched_log_likelihood(bs._h, B, N, pitch, Jr, Jc, p(b["coeffs"]), 5, p(T), p(D), p(Y),
                                        p(ld), p(q), p(ll), p(st), _capi.HOST_POINTERS)
    assert rc == 0 and np.all(st == 0)
    assert rel_sum(ld, ref[:, 0], N) < RTOL_SCALAR and rel_sum(q, ref[:, 1], N) < RTOL_SCALAR
    dev = torch.device("cuda:0")
    dT, dD, dY, dC = (torch.from_numpy(a).to(dev) for a in (T, D, Y, b["coeffs"]))
    out = torch.empty((3, B), dtype=torch.float64, device=dev)
    dst = torch.empty(B, dtype=torch.int32, device=dev)
    lib.cel_set_stream(bs._h, C.c_void_p(torch.cuda.current_stream(dev).cuda_stream or 1))
    v = lambda a: C.c_void_p(a.data_ptr())  # noqa: E731
    rc = lib.cel_batched_log_likelihood(bs._h, B, N, pitch, Jr, Jc, v(dC), 5, v(dT), v(dD), v(dY),
                                        v(out[0]), v(out[1]), v(out[2]), v(dst), 0)
    torch.cuda.synchronize()
    assert rc == 0 and int(dst.sum()) == 0
    assert rel_sum(out[0].cpu().numpy(), ref[:, 0], N) < RTOL_SCALAR
    assert rel_sum(out[1].cpu().numpy(), ref[:, 1], N) < RTOL_SCALAR


def test_argument_errors(bs):
    import ctypes as C
    from celerite_b200 import _capi
    lib = bs._lib
    a = np.zeros(16)
    p = C.c_void_p(a.ctypes.data)
    assert lib.cel_batched_log_likelihood(bs._h, 1, 0, 0, 0, 1, p, 5, p, p, p, p, p, p, p, 1) == _capi.INVALID_ARGUMENT
    assert lib.cel_batched_log_likelihood(bs._h, 2, 8, 4, 0, 1, p, 5, p, p, p, p, p, p, p, 1) == _capi.DIMENSION_MISMATCH
    assert lib.cel_batched_log_likelihood(bs._h, 2, 8, 8, 0, 49, p, 197, p, p, p, p, p, p, p, 1) == _capi.UNSUPPORTED   # J = 98 > 96
    assert lib.cel_batched_log_likelihood(bs._h, 0, 8, 8, 0, 1, p, 5, p, p, p, p, p, p, p, 1) == _capi.OK


@pytest.mark.parametrize("B,N,Jr,Jc,device", [(200, 700, 0, 1, True), (97, 403, 1, 1, False), (64, 300, 0, 2, True),
                                              (20, 500, 0, 4, True), (20, 500, 2, 3, False)])
def test_ragged_batch(bs, B, N, Jr, Jc, device):
    """Series of different lengths in one padded batch (cel_batched_log_likelihood_ragged): every
    series must get the log-likelihood of its own first lengths[b] points, whatever the padding
    holds — here zeros, NaN-free garbage and a series of length 0 and of length N."""
    from celerite_b200 import synth
    b = synth.make_batch(B, N, Jr, Jc, seed=B + N)
    rng = np.random.default_rng(B)
    lengths = rng.integers(1, N + 1, B).astype(np.int32)
    lengths[0], lengths[1], lengths[2] = N, 0, 1
    t, d, y = b["t"].copy(), b["diag"].copy(), b["y"].copy()
    for i in range(B):
        n = lengths[i]
        t[i, n:] = 0.0 if i % 2 else rng.uniform(-5, 5, N - n)      # padding: zeros or unsorted garbage
        d[i, n:] = rng.uniform(-1, 1, N - n)
        y[i, n:] = 1e6
    P = b["coeffs"].shape[1]
    chk = oracle.best()
    if device:
        import torch
        tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to("cuda:0")  # noqa: E731
        ll, logdet, quad, status = bs.log_likelihood(tt(b["coeffs"]), tt(t), tt(d), tt(y), Jr, Jc, return_parts=True,
                                                     lengths=tt(lengths))
        torch.cuda.synchronize()
        ll, logdet, quad, status = (x.cpu().numpy() for x in (ll, logdet, quad, status))
    else:
        ll, logdet, quad, status = bs.log_likelihood(b["coeffs"], t, d, y, Jr, Jc, return_parts=True, lengths=lengths)
    assert np.all(status == 0)
    for i in range(B):
        n = int(lengths[i])
        if n == 0:
            assert ll[i] == 0.0 and logdet[i] == 0.0 and quad[i] == 0.0
            continue
        ref = chk.batched_loglik(1, n, n, Jr, Jc, b["coeffs"][i:i + 1], P, b["t"][i:i + 1, :n].copy(),
                                 b["diag"][i:i + 1, :n].copy(), b["y"][i:i + 1, :n].copy(), nthreads=1)
        assert ref[0, 2] == 0
        assert rel_sum(logdet[i], ref[0, 0], n) < RTOL_SCALAR
        assert rel_sum(quad[i], ref[0, 1], n) < RTOL_SCALAR
        assert rel_sum(ll[i], -0.5 * (ref[0, 0] + ref[0, 1] + n * np.log(2 * np.pi)), n) < RTOL_SCALAR


def test_no_kernel_writes_behind_its_scratch():
    """CEL_DEBUG_GUARD: internal device buffers at exactly the requested size with a guard band behind each;
    tools/guard_check.py drives the gradient paths (two-kernel, spans, pieces, cooperative scan over a batch and
    over one long series), the term maps and the ragged compute at small sizes and asks
    cel_debug_guards_intact after every call.  A fresh process, since the variable is read once."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, CEL_DEBUG_GUARD="1")
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "guard_check.py")], cwd=root, env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "guards intact" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
