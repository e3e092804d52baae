"""GPU parity of the chunked two-pass scans (scan.cu): solve and dot_L on long series against
the CPU oracle, through the drop-in host API (cel_solve / cel_dot_L switch to the chunked path
at N >= 4096), the device API, and the multi-GPU span protocol emulated on one GPU."""
import numpy as np
import pytest

import oracle
from tests.util import RTOL_SOLVE, rel_inf

pytestmark = pytest.mark.gpu
E0, E2 = np.empty(0), np.empty((0, 0))


def _problem(N, Jr, Jc, seed):
    from celerite_b200 import synth
    b = synth.make_batch(1, N, Jr, Jc, seed=seed)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], Jr, Jc)
    return oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0]), b["y"][0]


@pytest.mark.parametrize("N,Jr,Jc,nrhs", [(4096, 0, 2, 1), (5003, 1, 1, 2), (16384, 0, 1, 1), (50001, 1, 1, 2), (40000, 0, 4, 3), (20000, 2, 6, 1),
                                          (17000, 0, 16, 1), (60160, 0, 1, 1), (20001, 1, 0, 2), (30000, 0, 2, 5), (33000, 0, 4, 8)])
def test_host_api_uses_chunked_path(N, Jr, Jc, nrhs):
    from celerite_b200 import solver
    p, y = _problem(N, Jr, Jc, seed=N)
    chk = oracle.best()
    f = chk.factor(p)
    s = solver.CholeskySolver()
    s.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, E0, E2, E2, p.x, p.diag)
    rhs = np.random.default_rng(N).standard_normal((N, nrhs))
    rhs[:, 0] = y
    n0 = s.launch_count()
    z = s.solve(rhs)
    assert s.launch_count() - n0 > 4          # the multi-kernel chunked path, not the 2-kernel serial one
    assert rel_inf(z, chk.solve(f, rhs)) < RTOL_SOLVE
    assert rel_inf(s.dot_L(rhs), chk.dot_L(f, rhs)) < RTOL_SOLVE
    # second call reuses the cached chunk maps
    assert rel_inf(s.solve(rhs), chk.solve(f, rhs)) < RTOL_SOLVE
    # y^T K^-1 y (cholesky.h:326-401): chunked forward sweep + ordered reduction
    n1 = s.launch_count()
    q, qref = s.dot_solve(y), chk.dot_solve(f, y)
    assert s.launch_count() - n1 > 2
    assert abs(q - qref) / max(abs(qref), N) < 1e-10


def test_device_api_and_span_protocol_on_one_gpu():
    """Three spans with warm-up halos on ONE GPU, the all-gather done by hand: must reproduce
    the single-span result (and the oracle)."""
    import torch
    from celerite_b200 import longseries as ls
    N, Jr, Jc = 60000, 0, 2
    p, y = _problem(N, Jr, Jc, seed=77)
    chk = oracle.best()
    f = chk.factor(p)
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(1)
    B = np.stack([y, rng.standard_normal(N)], axis=1)
    zref, lref = chk.solve(f, B), chk.dot_L(f, B)

    whole = ls.LongSeries(0)
    whole.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, p.x, p.diag)
    bt = torch.from_numpy(B).to(dev)
    assert rel_inf(whole.solve(bt).cpu().numpy(), zref) < RTOL_SOLVE
    assert rel_inf(whole.dot_L(bt).cpu().numpy(), lref) < RTOL_SOLVE
    assert rel_inf(whole.solve(bt[:, 0]).cpu().numpy(), zref[:, 0]) < RTOL_SOLVE

    world, halo = 3, 4000
    spans = [ls.span_bounds(N, world, r) for r in range(world)]
    locs = []
    for (lo, hi) in spans:
        h = min(halo, lo)
        L = ls.LongSeries(0)
        L.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp,
                  p.x[lo - h:hi], p.diag[lo - h:hi], halo=h)
        locs.append(L)

    def sweep(op, cols_per_rank):
        summ = torch.stack([L.summarize(op, c) for L, c in zip(locs, cols_per_rank)])
        outs = []
        for r, (L, c) in enumerate(zip(locs, cols_per_rank)):
            carry = ls.fold_carries(summ, r, op, c.shape[0], L.J)
            # the device-side fold (cel_span_fold) is the same arithmetic in one kernel
            dev_carry = L.fold(op, summ.contiguous(), c.shape[0], world, r)
            scale = max(1.0, float(carry.abs().max()))
            assert float((dev_carry - carry).abs().max()) / scale < 1e-12
            # summaries of all ranks were taken before any apply: pass 1 must be redone per rank
            outs.append(L.apply(op, c, carry, reuse_summary=False))
        return outs

    cols = [torch.from_numpy(np.ascontiguousarray(B[lo:hi].T)).to(dev) for (lo, hi) in spans]
    xd = sweep(ls.OP_SOLVE_FORWARD, cols)
    z = sweep(ls.OP_SOLVE_BACKWARD, xd)
    zz = np.concatenate([t.cpu().numpy().T for t in z], axis=0)
    assert rel_inf(zz, zref) < RTOL_SOLVE
    lz = sweep(ls.OP_DOT_L, cols)
    ll = np.concatenate([t.cpu().numpy().T for t in lz], axis=0)
    assert rel_inf(ll, lref) < RTOL_SOLVE
    # scalars: the spans' log-determinant partials and shares of y^T K^-1 y add up to the series'
    ld = sum(float(L.log_determinant(dev)) for L in locs)
    assert abs(ld - f.log_det) / N < 1e-12
    quad = sum(float(L.quad(x[0])) for L, x in zip(locs, xd))
    assert abs(quad - chk.dot_solve(f, y)) / N < 1e-10
    assert abs(float(whole.log_determinant(dev)) - f.log_det) / N < 1e-12
    llref = -0.5 * (chk.dot_solve(f, y) + f.log_det + N * np.log(2 * np.pi))
    assert abs(float(whole.log_likelihood(bt[:, 0].contiguous())) - llref) / N < 1e-10


def test_full_size_properties_cfg4_like():
    """N = 2e6, J = 8 (a fifth of config 4): size-independent properties.
    K (K^-1 b) = b through the stateless dot() + measurement variances, and linearity."""
    import torch
    from celerite_b200 import longseries as ls, solver
    N, Jr, Jc = 2_000_000, 0, 4
    p, y = _problem(N, Jr, Jc, seed=5)
    dev = torch.device("cuda:0")
    L = ls.LongSeries(0)
    L.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, p.x, p.diag)
    b1 = torch.from_numpy(y).to(dev)
    b2 = torch.from_numpy(np.random.default_rng(2).standard_normal(N)).to(dev)
    z1, z2, z12 = L.solve(b1), L.solve(b2), L.solve(b1 + 2.0 * b2)
    scale = float(z12.abs().max())
    assert float((z12 - (z1 + 2.0 * z2)).abs().max()) / scale < 1e-11
    s = solver.CholeskySolver()
    z = z1.cpu().numpy()
    kz = s.dot(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, E0, E2, E2, p.x, z)[:, 0]
    kz += p.diag * z
    assert rel_inf(kz, y) < 1e-8


def _factor_diff(state, f):
    comp, n, j, ld, phi, u, W, D = state
    scale = lambda a: max(1e-300, float(np.max(np.abs(a))))  # noqa: E731
    return max(np.max(np.abs(D - f.D)) / scale(f.D), np.max(np.abs(W.T - f.W)) / scale(f.W),
               np.max(np.abs(phi.T - f.phi)) / scale(f.phi), np.max(np.abs(u.T - f.u)) / scale(f.u))


@pytest.mark.parametrize("N,Jr,Jc,segmented", [(200000, 0, 4, True), (70001, 1, 1, True), (150000, 0, 1, True), (20000, 0, 2, None),
                                               (100003, 2, 6, None)])
def test_segmented_compute_matches_serial_factor(N, Jr, Jc, segmented):
    """cel_compute on a long series factorises time segments side by side behind warm-up halos
    (capi_single.cu: plan_segments).  The stored factor and log-determinant must be the serial ones."""
    from celerite_b200 import solver
    p, y = _problem(N, Jr, Jc, seed=N + 1)
    chk = oracle.best()
    f = chk.factor(p)
    s = solver.CholeskySolver()
    n0 = s.launch_count()
    s.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, E0, E2, E2, p.x, p.diag)
    # segment sweep + finish; the serial path is one launch (segmented=None: the planner decides)
    assert s.launch_count() - n0 == (2 if segmented else s.launch_count() - n0)
    assert abs(s.log_determinant() - f.log_det) / N < 1e-12
    assert _factor_diff(s.__getstate__(), f) < 1e-11
    assert rel_inf(s.solve(y), chk.solve(f, y)) < RTOL_SOLVE


def test_segmented_compute_falls_back_for_slow_kernels():
    """A term that barely decays over the whole series leaves nothing to cut: serial sweep."""
    from celerite_b200 import solver
    N = 80000
    p, y = _problem(N, 1, 1, seed=3)
    span = p.x[-1] - p.x[0]
    c_real = np.array(p.c_real, copy=True)
    c_real[0] = 0.5 / span                      # a real term with a > 0 stays positive definite
    pp = oracle.Problem(p.jitter, p.a_real, c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, p.x, p.diag)
    chk = oracle.best()
    f = chk.factor(pp)
    s = solver.CholeskySolver()
    n0 = s.launch_count()
    s.compute(pp.jitter, pp.a_real, pp.c_real, pp.a_comp, pp.b_comp, pp.c_comp, pp.d_comp, E0, E2, E2, pp.x, pp.diag)
    assert s.launch_count() - n0 == 1
    assert _factor_diff(s.__getstate__(), f) < 1e-11


def test_segmented_compute_reports_the_first_negative_pivot():
    """cholesky.h:176 — the failure index must be the serial one even when it lies deep in a segment."""
    from celerite_b200 import solver
    N = 120000
    p, y = _problem(N, 0, 2, seed=9)
    diag = np.array(p.diag, copy=True)
    bad = 77777
    diag[bad] = -50.0 * (np.sum(p.a_comp) + 1.0)
    diag[bad + 20000] = diag[bad]
    s = solver.CholeskySolver()
    with pytest.raises(Exception) as ei:
        s.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, E0, E2, E2, p.x, diag)
    assert "n = %d" % bad in str(ei.value)
    assert not s.computed()


def test_segmented_compute_rejects_halos_that_did_not_converge(monkeypatch):
    """With the halos cut to 2 % of the bound the segments start from a wrong state; the check of
    every halo against its owner must notice and the series is redone serially (3 launches)."""
    from celerite_b200 import solver
    N = 150000
    p, y = _problem(N, 0, 2, seed=21)
    chk = oracle.best()
    f = chk.factor(p)
    monkeypatch.setenv("CEL_SEGMENT_HALO_SCALE", "0.02")
    s = solver.CholeskySolver()
    n0 = s.launch_count()
    s.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, E0, E2, E2, p.x, p.diag)
    assert s.launch_count() - n0 == 3
    assert abs(s.log_determinant() - f.log_det) / N < 1e-12
    assert _factor_diff(s.__getstate__(), f) < 1e-11


def _grad_case(N, Jr, Jc, seed):
    from celerite_b200 import solver
    from tests.util import RTOL_GRAD, RTOL_SCALAR, rel_grad, rel_sum
    p, y = _problem(N, Jr, Jc, seed=seed)
    st, vref, gref = oracle.best().grad(p, y)
    assert st == 0
    s = solver.CholeskySolver()
    n0 = s.launch_count()
    val, g = s.grad_log_likelihood(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp,
                                   E0, E2, E2, p.x, y, p.diag)
    assert rel_sum(val, vref, N) < RTOL_SCALAR
    assert rel_grad(g, gref) < RTOL_GRAD, (g, gref)
    return s.launch_count() - n0


@pytest.mark.parametrize("N,Jr,Jc", [(150000, 0, 2), (100001, 1, 1), (200000, 0, 4)])
def test_span_parallel_gradient_matches_reference(N, Jr, Jc):
    """grad_log_likelihood of one long series runs as time spans side by side with the adjoint as an exact
    two-pass scan (grad.cu, scan modes: probe, fold, final, finish): value and gradient against the
    reference's automatic differentiation."""
    assert _grad_case(N, Jr, Jc, seed=N + 7) == 4


def test_gradient_just_above_the_span_threshold():
    """N = 20000: whichever path the planner picks must meet the bar."""
    assert _grad_case(20000, 0, 2, seed=77) in (1, 4, 5, 9)


def test_span_parallel_gradient_rejects_short_halos(monkeypatch):
    """Warm-up halos cut to 2 % of the bound: neighbouring spans disagree at their common edge (W, D, f of
    the same step), the check notices, and the second attempt -- ONE forward sweep over the series for the
    checkpoints, the adjoint scan as before -- gives the answer (4 + 5 launches)."""
    monkeypatch.setenv("CEL_SEGMENT_HALO_SCALE", "0.02")
    assert _grad_case(120000, 0, 2, seed=31) == 9


@pytest.mark.parametrize("N,Jr,Jc", [(90000, 0, 2), (70001, 2, 1), (50000, 0, 8), (40000, 0, 13)])
def test_adjoint_scan_behind_a_serial_forward_sweep(monkeypatch, N, Jr, Jc):
    """The path for kernels that forget too slowly for warm-up halos, forced: forward sweep of the whole
    series (mode 3), then probe, fold, final, finish.  No halo anywhere, so nothing is approximate."""
    monkeypatch.setenv("CEL_GRAD_SCAN_SERIAL_FORWARD", "1")
    assert _grad_case(N, Jr, Jc, seed=N + 3) == 5


def test_halo_and_cool_down_spans_of_round_one_still_agree(monkeypatch):
    """CEL_GRAD_SPANS_V1: the span sweep with warm-up and cool-down halos (kept as an A/B knob)."""
    monkeypatch.setenv("CEL_GRAD_SPANS_V1", "1")
    assert _grad_case(150000, 0, 2, seed=150007) == 2


@pytest.mark.parametrize("N,Jr,Jc,nrhs", [(50001, 1, 2, 2), (4096, 0, 1, 1), (40000, 2, 0, 3), (70000, 0, 8, 1)])
def test_stateless_dot_on_long_series(N, Jr, Jc, nrhs):
    """K z (cholesky.h:444-590) from N = 4096 on runs as chunks side by side with closed-form
    carries (sweeps.cu: k_dot_chunk)."""
    from celerite_b200 import solver
    p, y = _problem(N, Jr, Jc, seed=N + 3)
    z = np.random.default_rng(N).standard_normal((N, nrhs))
    s = solver.CholeskySolver()
    n0 = s.launch_count()
    got = s.dot(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, E0, E2, E2, p.x, z)
    assert s.launch_count() - n0 == 3
    assert rel_inf(got, oracle.best().dot(p, z)) < RTOL_SOLVE


@pytest.mark.parametrize("N,Jr,Jc,M", [(40000, 1, 2, 25000), (4097, 0, 1, 7), (60001, 0, 4, 90000)])
def test_predict_on_long_series(N, Jr, Jc, M):
    """predict (cholesky.h:599-698) from N = 4096 on: chunked solve, then the two merge sweeps by
    chunks of data points.  Test points before the first and after the last datum, on data times,
    and repeated."""
    from celerite_b200 import solver
    p, y = _problem(N, Jr, Jc, seed=N + 5)
    chk = oracle.best()
    f = chk.factor(p)
    rng = np.random.default_rng(M)
    lo, hi = p.x[0], p.x[-1]
    xs = rng.uniform(lo - 0.05 * (hi - lo), hi + 0.05 * (hi - lo), M)
    k = min(M // 3, 1000)
    xs[:k] = p.x[rng.integers(0, N, k)]                 # exactly on data times
    if M > 4:
        xs[-2] = xs[-1]                                  # a repeated test point
    xs = np.sort(xs)
    s = solver.CholeskySolver()
    s.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, E0, E2, E2, p.x, p.diag)
    assert rel_inf(s.predict(y, xs), chk.predict(p, f, y, xs)) < RTOL_SOLVE


def test_mgpu_entry_points_with_a_world_of_one():
    """cel_comm_init / cel_compute_mgpu / cel_solve_mgpu / cel_dot_L_mgpu / cel_log_likelihood_mgpu through
    the C ABI with world = 1 (no NCCL needed: the collectives degenerate, the rest of the path is the one
    every rank of a larger world runs) against the reference.  tests/test_gpu_multi.py does two GPUs."""
    import ctypes as C
    import torch
    from celerite_b200 import _capi
    N, Jr, Jc = 50_000, 1, 2
    p, y = _problem(N, Jr, Jc, seed=123)
    chk = oracle.best()
    f = chk.factor(p)
    lib = _capi.lib()
    h = _capi.create(0)
    try:
        dev = torch.device("cuda:0")
        ptr = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
        lo, hi = C.c_int64(), C.c_int64()
        assert lib.cel_mgpu_bounds(N, 1, 0, C.byref(lo), C.byref(hi)) == _capi.OK and (lo.value, hi.value) == (0, N)
        _capi.check(lib.cel_comm_init(h, 1, 0, None), h)
        ar, cr, ac, bc, cc, dc = (np.ascontiguousarray(v, dtype=np.float64) for v in
                                  (p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp))
        _capi.check(lib.cel_compute_mgpu(h, float(p.jitter), Jr, ptr(ar), ptr(cr), Jc, ptr(ac), ptr(bc), ptr(cc), ptr(dc),
                                         N, ptr(p.x), ptr(p.diag), 0), h)
        rhs = np.stack([y, np.random.default_rng(5).standard_normal(N)], axis=0)      # two columns, column-major N x 2
        d_in = torch.from_numpy(rhs).to(dev)
        d_out = torch.empty_like(d_in)
        v = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        lib.cel_set_stream(h, C.c_void_p(torch.cuda.current_stream(dev).cuda_stream or 1))
        _capi.check(lib.cel_solve_mgpu(h, 2, v(d_in), v(d_out)), h)
        torch.cuda.synchronize(dev)
        assert rel_inf(d_out.cpu().numpy().T, chk.solve(f, rhs.T.copy())) < RTOL_SOLVE
        _capi.check(lib.cel_dot_L_mgpu(h, 2, v(d_in), v(d_out)), h)
        torch.cuda.synchronize(dev)
        assert rel_inf(d_out.cpu().numpy().T, chk.dot_L(f, rhs.T.copy())) < RTOL_SOLVE
        out3 = (C.c_double * 3)()
        _capi.check(lib.cel_log_likelihood_mgpu(h, v(d_in[0].contiguous()), out3), h)
        quad = chk.dot_solve(f, y)
        ll = -0.5 * (quad + f.log_det + N * np.log(2 * np.pi))
        assert abs(out3[0] - f.log_det) / abs(f.log_det) < 1e-10
        assert abs(out3[1] - quad) / abs(quad) < 1e-10
        assert abs(out3[2] - ll) / abs(ll) < 1e-10
        # a sweep before cel_compute_mgpu is refused, not run
        h2 = _capi.create(0)
        assert lib.cel_solve_mgpu(h2, 1, v(d_in), v(d_out)) == _capi.NOT_COMPUTED
        lib.cel_destroy(h2)
    finally:
        lib.cel_destroy(h)
        if hasattr(chk, "release"):
            chk.release(f)


def test_high_q_oscillator_takes_the_serial_forward_sweep_by_itself():
    """Two SHO terms of quality factor 300 and 1,500 (c = w0 / 2Q ~ 1e-3): the factorisation forgets over
    ~4e4 time units, far longer than the series, so no warm-up halo can be planned; the gradient must come
    out of the forward sweep + adjoint scan (5 launches) and match the reference's AD."""
    from celerite_b200 import solver, synth
    from tests.util import RTOL_GRAD, RTOL_SCALAR, rel_grad, rel_sum
    N = 60000
    b = synth.make_batch(1, N, 0, 2, seed=5)
    a0, b0, c0, d0 = synth.sho_coefficients(1.3, 300.0, 0.9)
    a1, b1, c1, d1 = synth.sho_coefficients(0.6, 1500.0, 4.0)
    p = oracle.Problem(synth.JITTER, E0, E0, np.array([a0, a1]), np.array([b0, b1]), np.array([c0, c1]),
                       np.array([d0, d1]), b["t"][0], b["diag"][0])
    y = b["y"][0]
    st, vref, gref = oracle.best().grad(p, y)
    assert st == 0
    s = solver.CholeskySolver()
    n0 = s.launch_count()
    val, g = s.grad_log_likelihood(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp,
                                   E0, E2, E2, p.x, y, p.diag)
    assert s.launch_count() - n0 == 5
    assert rel_sum(val, vref, N) < RTOL_SCALAR
    assert rel_grad(g, gref) < RTOL_GRAD, (g, gref)


@pytest.mark.parametrize("N,Jr,Jc", [(1024, 0, 1), (1500, 1, 2), (4096, 1, 1), (6000, 0, 13)])
def test_short_series_gradient_is_a_forward_sweep_and_an_adjoint_scan(N, Jr, Jc):
    """From N = 1,024 on the single-series gradient runs as ONE forward sweep (mode 3) and the exact adjoint scan
    over spans of >= 256 steps (probe, fold, final, finish: 5 launches); N = 4,096 with one real and one complex
    term is BASELINE config 1."""
    assert _grad_case(N, Jr, Jc, seed=N + 11) == 5


def test_series_below_the_scan_threshold_take_the_serial_sweep():
    assert _grad_case(1000, 1, 1, seed=5) == 1
