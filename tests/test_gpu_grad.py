"""GPU parity of the reverse-mode gradient (K6) against the reference's forward-mode AD gradient
(celerite/solver.cpp:347-463): committed golden vectors, and oracle/_ref or the C port on fresh
seeded inputs.  Tolerance 1e-8 relative per component (north_star)."""
import numpy as np
import pytest

import oracle
from tests.util import RTOL_GRAD, RTOL_SCALAR, golden_names, rel_grad, rel_sum

pytestmark = pytest.mark.gpu
NAMES = golden_names()
E0, E2 = np.empty(0), np.empty((0, 0))


@pytest.fixture(scope="module")
def S():
    from celerite_b200 import solver
    return solver


def _grad(S, p, y):
    s = S.CholeskySolver()
    return s.grad_log_likelihood(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp,
                                 p.A if p.A is not None else E0, p.U if p.U is not None else E2,
                                 p.V if p.V is not None else E2, p.x, y, p.diag)


@pytest.mark.parametrize("name", NAMES)
def test_golden_grad(golden, S, name):
    c = golden[name]
    p = c.problem
    if c.status != 0:
        with pytest.raises(S.LinAlgError):
            _grad(S, p, c.y)
        return
    val, g = _grad(S, p, c.y)
    assert g.shape == (p.P,)
    assert rel_sum(val, c.grad_val, p.N) < RTOL_SCALAR          # incl. the pi*log(N) constant
    assert rel_grad(g, c.grad) < RTOL_GRAD, (g, c.grad)


def test_jitter_quirk(S):
    """solver.cpp:377-389: jitter <= DBL_EPSILON forces slot 0 to exactly 0.0 (slot still present)."""
    rng = np.random.default_rng(0)
    N = 64
    x = np.sort(rng.uniform(0, 5, N)); diag = rng.uniform(0.1, 0.3, N) ** 2; y = np.sin(x)
    s = S.CholeskySolver()
    val, g = s.grad_log_likelihood(0.0, [1.0], [0.5], [], [], [], [], E0, E2, E2, x, y, diag)
    assert g.shape == (3,) and g[0] == 0.0 and g[1] != 0.0
    val, g = s.grad_log_likelihood(1e-3, [1.0], [0.5], [], [], [], [], E0, E2, E2, x, y, diag)
    assert g[0] != 0.0


@pytest.mark.parametrize("N,Jr,Jc", [(4096, 1, 1), (2000, 0, 1), (1000, 0, 2), (1537, 2, 3), (3000, 0, 8), (900, 0, 16)])
def test_random_vs_oracle(S, N, Jr, Jc):
    from celerite_b200 import synth
    chk = oracle.best()
    b = synth.make_batch(1, N, Jr, Jc, seed=300 + N)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], Jr, Jc)
    p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0])
    st, vref, gref = chk.grad(p, b["y"][0])
    assert st == 0
    val, g = _grad(S, p, b["y"][0])
    assert rel_sum(val, vref, N) < RTOL_SCALAR
    assert rel_grad(g, gref) < RTOL_GRAD, (g, gref)


@pytest.mark.parametrize("B,N,Jr,Jc", [(70, 500, 0, 1), (33, 700, 1, 1), (9, 400, 0, 4), (3, 300, 0, 16)])
def test_batched_grad(B, N, Jr, Jc):
    import ctypes as C
    from celerite_b200 import _capi, synth
    from celerite_b200.batched import BatchedSolver
    b = synth.make_batch(B, N, Jr, Jc, seed=B + N)
    P = b["coeffs"].shape[1]
    bs = BatchedSolver()
    val, g, status = bs.grad_log_likelihood(b["coeffs"], b["t"], b["diag"], b["y"], Jr, Jc)
    assert np.all(status == 0)
    chk = oracle.best()
    for i in range(B):
        j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][i], Jr, Jc)
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][i], b["diag"][i])
        st, vref, gref = chk.grad(p, b["y"][i])
        vref += 0.5 * np.pi * np.log(N)   # batched value carries no constant
        assert rel_sum(val[i], vref, N) < RTOL_SCALAR
        assert rel_grad(g[i], gref) < RTOL_GRAD


@pytest.mark.parametrize("B,N,Jc", [(100, 500, 1), (33, 2000, 1), (70, 403, 2), (5, 1000, 2), (40, 17, 1), (3, 1, 1),
                                    (40, 20000, 1), (300, 4098, 1), (2, 262, 1)])
def test_batched_grad_lockstep_device(B, N, Jc):
    """The thread-per-series gradient kernel (all-complex J = 2, 4; device-resident, TMA path) must
    agree with the reference gradient and with the cooperative kernel."""
    import os
    import torch
    from celerite_b200 import synth
    from celerite_b200.batched import BatchedSolver
    b = synth.make_batch(B, N, 0, Jc, seed=11 * B + N)
    dev = torch.device("cuda:0")
    tt = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    bs = BatchedSolver()
    val, g, status = bs.grad_log_likelihood(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"]), 0, Jc)
    torch.cuda.synchronize()
    val, g, status = val.cpu().numpy(), g.cpu().numpy(), status.cpu().numpy()
    assert np.all(status == 0)
    chk = oracle.best()
    for i in sorted(set(list(range(0, B, max(1, B // 9))) + [B - 1])):
        j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][i], 0, Jc)
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][i], b["diag"][i])
        st, vref, gref = chk.grad(p, b["y"][i])
        vref += 0.5 * np.pi * np.log(N)
        assert rel_sum(val[i], vref, N) < RTOL_SCALAR
        assert rel_grad(g[i], gref) < RTOL_GRAD, (i, g[i], gref)


@pytest.mark.parametrize("B,N,Jr,Jc", [(40, 400, 0, 1), (21, 350, 1, 1), (12, 300, 0, 4)])
def test_ragged_batched_grad(B, N, Jr, Jc):
    """cel_batched_grad_log_likelihood_ragged: every series is differentiated over its own length,
    whatever its padding holds (incl. a series of length 0, of length 1 and of length N)."""
    import torch
    from celerite_b200 import synth
    from celerite_b200.batched import BatchedSolver
    b = synth.make_batch(B, N, Jr, Jc, seed=5 * B + N)
    rng = np.random.default_rng(B)
    lengths = rng.integers(2, N + 1, B).astype(np.int32)
    lengths[0], lengths[1], lengths[2] = N, 0, 1
    t, d, y = b["t"].copy(), b["diag"].copy(), b["y"].copy()
    for i in range(B):
        n = lengths[i]
        t[i, n:] = rng.uniform(-5, 5, N - n); d[i, n:] = -1.0; y[i, n:] = 1e6
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to("cuda:0")  # noqa: E731
    val, g, status = BatchedSolver().grad_log_likelihood_ragged(tt(b["coeffs"]), tt(t), tt(d), tt(y), tt(lengths), Jr, Jc)
    torch.cuda.synchronize()
    val, g, status = val.cpu().numpy(), g.cpu().numpy(), status.cpu().numpy()
    assert np.all(status == 0)
    chk = oracle.best()
    for i in range(B):
        n = int(lengths[i])
        if n == 0:
            assert val[i] == 0.0 and np.all(g[i] == 0.0)
            continue
        j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][i], Jr, Jc)
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][i, :n].copy(), b["diag"][i, :n].copy())
        st, vref, gref = chk.grad(p, b["y"][i, :n].copy())
        vref += 0.5 * np.pi * np.log(n)   # batched value carries no constant
        assert rel_sum(val[i], vref, n) < RTOL_SCALAR
        assert rel_grad(g[i], gref) < RTOL_GRAD, (i, n, g[i], gref)


@pytest.mark.parametrize("Jc", [1, 2])
def test_two_kernel_gradient_redoes_series_that_leave_the_fast_path(Jc):
    """J = 2 and J = 4 batches take the two-kernel gradient (lockstep_grad2.cu).  Series its forward kernel flags
    -- pivots outside [2^-500, 2^500], a negative pivot, |d t| >= 1e9 -- are redone by the first-generation
    kernel: every series must still match the reference's gradient (solver.cpp:347-463), failures must
    come out as status 3 with NaNs, ordinary neighbours in the same warp must stay correct."""
    import torch
    from celerite_b200 import synth
    from celerite_b200.batched import BatchedSolver
    B, N = 150, 403
    b = synth.make_batch(B, N, 0, Jc, seed=77)
    huge, neg, far, nojit = 5, 40, 97, 130
    # series `huge`: K -> 1e200 K and y -> 1e100 y, so that every pivot is ~1e200 (outside the fast path's range).
    # The reference's forward-mode derivatives overflow there, so it is checked through the scaling law:
    # value - N/2 ln s, gradient in (jitter, a, b) times s, in (c, d) unchanged, against the unscaled series.
    s2 = 1e200
    unscaled = {k: b[k][huge].copy() for k in ("coeffs", "diag", "y")}
    b["coeffs"][huge, :1 + 2 * Jc] *= s2                     # jitter, every a and every b
    b["diag"][huge] *= s2; b["y"][huge] *= 1e100
    b["diag"][neg, 200] = -50.0                              # a negative pivot half-way
    b["t"][far] += 3.0e9                                     # |d t| beyond the branch-free sincos
    b["coeffs"][nojit, 0] = 0.0                              # jitter <= eps: slot 0 is forced to 0
    dev = torch.device("cuda:0")
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    bs = BatchedSolver()
    val, g, status = bs.grad_log_likelihood(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"]), 0, Jc)
    val, g, status = val.cpu().numpy(), g.cpu().numpy(), status.cpu().numpy()
    chk = oracle.best()
    for i in range(B):
        src = unscaled if i == huge else {k: b[k][i] for k in ("coeffs", "diag", "y")}
        j, ar, cr, ac, bc, cc, dc = synth.split_row(src["coeffs"], 0, Jc)
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][i], src["diag"])
        st, vref, gref = chk.grad(p, src["y"])
        if i == neg:
            assert st != 0 and status[i] == 3 and np.isnan(val[i]) and np.all(np.isnan(g[i]))
            continue
        if i == huge:
            vref -= 0.5 * N * np.log(s2)
            # (the reference's gradient of the unscaled series IS the comparison value once the scaled
            #  components -- jitter, a, b -- of the GPU gradient are multiplied by s)
            g[i, :1 + 2 * Jc] *= s2
        assert st == 0 and status[i] == 0, (i, st, status[i])
        vref += 0.5 * np.pi * np.log(N)
        tol_s, tol_g = (1e-9, 1e-6) if i == far else (RTOL_SCALAR, RTOL_GRAD)   # d t ~ 1e10: argument rounding
        assert abs(val[i] - vref) <= tol_s * max(abs(vref), N), (i, val[i], vref)
        if i == far:
            # d/d(d) is a sum of terms t_n * (...) with t_n ~ 3e9 that cancels to ~1e2: its rounding error is
            # ~1e9 times that of the other components, in the reference's forward-mode sum as much as here
            nd = 1 + 3 * Jc                      # the d components are the last Jc
            assert rel_grad(g[i, :nd], gref[:nd]) < tol_g, (i, g[i], gref)
            assert np.all(np.abs(g[i, nd:] - gref[nd:]) < 1e-3 * np.max(np.abs(gref[nd:]))), (i, g[i], gref)
            continue
        assert rel_grad(g[i], gref) < tol_g, (i, g[i], gref)
    assert g[nojit, 0] == 0.0 and g[0, 0] != 0.0


@pytest.mark.parametrize("B,N,Jr,Jc", [(40, 30000, 1, 1), (100, 8192, 0, 3), (17, 20001, 2, 0), (6, 50000, 0, 8)])
def test_few_long_series_take_the_batched_adjoint_scan(B, N, Jr, Jc):
    """Batches that leave most of the GPU empty with a group of lanes per whole series (few series, N >= 4096,
    J <= 16) run as a forward sweep per series followed by the exact adjoint scan over spans of every series
    (grad.cu, launch_grad_batch_scan: 6 launches); value, gradient and status against the reference, with one
    series that is not positive definite."""
    import torch
    from celerite_b200 import synth
    from celerite_b200.batched import BatchedSolver
    b = synth.make_batch(B, N, Jr, Jc, seed=B + N)
    bad = B // 2
    b["diag"][bad, N // 3] = -80.0
    dev = torch.device("cuda:0")
    tt = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    bs = BatchedSolver()
    n0 = bs._lib.cel_launch_count(bs._h)
    val, g, status = bs.grad_log_likelihood(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"]), Jr, Jc)
    torch.cuda.synchronize()
    assert bs._lib.cel_launch_count(bs._h) - n0 == 6
    val, g, status = val.cpu().numpy(), g.cpu().numpy(), status.cpu().numpy()
    chk = oracle.best()
    for i in sorted(set([0, 1, bad, B - 1] + list(range(0, B, max(1, B // 5))))):
        j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][i], Jr, Jc)
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][i], b["diag"][i])
        st, vref, gref = chk.grad(p, b["y"][i])
        if i == bad:
            assert st != 0 and status[i] == 3 and np.isnan(val[i]) and np.all(np.isnan(g[i]))
            continue
        assert st == 0 and status[i] == 0
        vref += 0.5 * np.pi * np.log(N)
        assert rel_sum(val[i], vref, N) < RTOL_SCALAR
        assert rel_grad(g[i], gref) < RTOL_GRAD, (i, g[i], gref)


def test_two_kernel_gradient_in_pieces(monkeypatch):
    """The J = 2 gradient goes through a large batch in pieces that share one scratch area (checkpoints,
    flags, exchange rows); forced here to pieces of 448 series, the last one shorter and -- being small --
    swept in spans where the full pieces are not.  Every series against the reference."""
    import torch
    from celerite_b200 import synth
    from celerite_b200.batched import BatchedSolver
    monkeypatch.setenv("CEL_GRAD2_PIECE", "448")
    B, N = 1000, 1200
    b = synth.make_batch(B, N, 0, 1, seed=99)
    dev = torch.device("cuda:0")
    tt = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    bs = BatchedSolver()
    val, g, status = bs.grad_log_likelihood(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"]), 0, 1)
    torch.cuda.synchronize()
    val, g, status = val.cpu().numpy(), g.cpu().numpy(), status.cpu().numpy()
    assert np.all(status == 0)
    rv, rg, rs = oracle.best().batched_grad(B, N, N, 0, 1, b["coeffs"], b["coeffs"].shape[1], b["t"], b["diag"], b["y"], 8)
    rv = rv + 0.5 * np.pi * np.log(N)
    assert np.max(np.abs(val - rv) / np.maximum(np.abs(rv), 1.0)) < RTOL_SCALAR
    assert max(rel_grad(g[i], rg[i]) for i in range(B)) < RTOL_GRAD


@pytest.mark.parametrize("B,N,Jc,device", [(100, 600, 1, True), (100, 600, 1, False), (40, 4096, 1, True), (12, 500, 2, True)])
def test_batched_grad_with_one_coefficient_row_for_all(B, N, Jc, device):
    """coeff_stride = 0: one kernel for every series of the batch (one MCMC walker, many light curves), through the
    two-kernel gradient (J = 2; whole series and, for (40, 4096), spans) and the one-kernel form (J = 4)."""
    import torch
    from celerite_b200 import synth
    from celerite_b200.batched import BatchedSolver
    b = synth.make_batch(B, N, 0, Jc, seed=B + N + Jc)
    row = np.ascontiguousarray(b["coeffs"][3])
    bs = BatchedSolver()
    if device:
        dev = torch.device("cuda:0")
        tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
        val, g, status = bs.grad_log_likelihood(tt(row), tt(b["t"]), tt(b["diag"]), tt(b["y"]), 0, Jc)
        torch.cuda.synchronize()
        val, g, status = val.cpu().numpy(), g.cpu().numpy(), status.cpu().numpy()
    else:
        val, g, status = bs.grad_log_likelihood(row, b["t"], b["diag"], b["y"], 0, Jc)
    assert np.all(status == 0) and val.shape == (B,) and g.shape == (B, 1 + 4 * Jc)
    chk = oracle.best()
    j, ar, cr, ac, bc, cc, dc = synth.split_row(row, 0, Jc)
    for i in sorted(set(list(range(0, B, max(1, B // 6))) + [B - 1])):
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][i], b["diag"][i])
        st, vref, gref = chk.grad(p, b["y"][i])
        vref += 0.5 * np.pi * np.log(N)
        assert rel_sum(val[i], vref, N) < RTOL_SCALAR
        assert rel_grad(g[i], gref) < RTOL_GRAD, (i, g[i], gref)


@pytest.mark.parametrize("B,N,Jr,Jc", [(2100, 300, 1, 1), (2500, 202, 2, 1), (2200, 150, 1, 0), (2300, 100, 2, 0),
                                       (2050, 90, 3, 0), (2100, 124, 4, 0)])
def test_batched_grad_with_real_terms_thread_per_series(B, N, Jr, Jc):
    """Real + complex terms with J <= 4 (RealTerm + ComplexTerm is BASELINE config 1's kernel) in batches of
    thousands take the two-kernel thread-per-series gradient over arrays (lockstep_grad2.cu, launch_grad_mixed:
    forward with checkpoints, reverse, and a redo pass that leaves at once): 3 launches, every checked series
    against the reference's AD; from host arrays as well."""
    import torch
    from celerite_b200 import synth
    from celerite_b200.batched import BatchedSolver
    b = synth.make_batch(B, N, Jr, Jc, seed=3 * B + N)
    dev = torch.device("cuda:0")
    tt = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    bs = BatchedSolver()
    n0 = bs._lib.cel_launch_count(bs._h)
    val, g, status = bs.grad_log_likelihood(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"]), Jr, Jc)
    torch.cuda.synchronize()
    assert bs._lib.cel_launch_count(bs._h) - n0 == 3
    val, g, status = val.cpu().numpy(), g.cpu().numpy(), status.cpu().numpy()
    hv, hg, hs = bs.grad_log_likelihood(b["coeffs"], b["t"], b["diag"], b["y"], Jr, Jc)
    assert np.all(status == 0) and np.all(hs == 0)
    assert np.allclose(hv, val, rtol=1e-13) and np.allclose(hg, g, rtol=1e-10, atol=1e-10)
    chk = oracle.best()
    for i in sorted(set(list(range(0, B, B // 7)) + [B - 1])):
        j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][i], Jr, Jc)
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][i], b["diag"][i])
        st, vref, gref = chk.grad(p, b["y"][i])
        vref += 0.5 * np.pi * np.log(N)
        assert rel_sum(val[i], vref, N) < RTOL_SCALAR
        assert rel_grad(g[i], gref) < RTOL_GRAD, (i, g[i], gref)


def test_batched_grad_with_real_terms_redoes_the_batch_when_a_series_leaves_the_fast_path():
    """One series with a negative pivot among 2,100: the forward kernel flags it, the cooperative kernel redoes the
    batch (guarded arithmetic, the reference's failure semantics): status 3 and NaN there, everything else right."""
    import torch
    from celerite_b200 import synth
    from celerite_b200.batched import BatchedSolver
    B, N, Jr, Jc = 2100, 160, 1, 1
    b = synth.make_batch(B, N, Jr, Jc, seed=4)
    bad = 1234
    b["diag"][bad, 77] = -60.0
    dev = torch.device("cuda:0")
    tt = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    bs = BatchedSolver()
    val, g, status = bs.grad_log_likelihood(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"]), Jr, Jc)
    val, g, status = val.cpu().numpy(), g.cpu().numpy(), status.cpu().numpy()
    assert status[bad] == 3 and np.isnan(val[bad]) and np.all(np.isnan(g[bad]))
    assert status.sum() == 3
    chk = oracle.best()
    for i in (0, bad - 1, bad + 1, B - 1):
        j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][i], Jr, Jc)
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][i], b["diag"][i])
        st, vref, gref = chk.grad(p, b["y"][i])
        vref += 0.5 * np.pi * np.log(N)
        assert rel_sum(val[i], vref, N) < RTOL_SCALAR
        assert rel_grad(g[i], gref) < RTOL_GRAD, (i, g[i], gref)
