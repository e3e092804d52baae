"""GPU parity at the REAL sizes of BASELINE.json's configs against the compiled reference (oracle.best()),
with PURE relative errors for scalars and norm-wise errors for vectors.  Next to every GPU-vs-reference
figure the reference's own round-off floor is printed (pytest -s): the same quantity from the other CPU
build of the same algorithm (the reference with FMA contraction against the plain-C restatement without).

    cfg2   4,096 series x N=2,000, J=2            log-likelihood, every series
    cfg5   8,192 series x N=1,000, J=4            log-likelihood, every series
    cfg3   one series N=100,000 at J=8 and J=32   value + gradient against the reference's forward-mode AD
                                                  (solver.cpp:347-463; ~22 s of CPU time at J=32)
    cfg4   one series N=10^7, J=8                 compute, solve, dot_L, log-likelihood over the WHOLE series
"""
import numpy as np
import pytest

import oracle
from tests.util import RTOL_GRAD, RTOL_SCALAR, RTOL_SOLVE, rel_grad, rel_inf, scalar_parity

pytestmark = pytest.mark.gpu
E0, E2 = np.empty(0), np.empty((0, 0))


@pytest.mark.parametrize("name,B,N,Jr,Jc", [("cfg2", 4096, 2000, 0, 1), ("cfg5", 8192, 1000, 0, 2)])
def test_batched_loglik_config_sized(name, B, N, Jr, Jc):
    from celerite_b200 import synth
    from celerite_b200.batched import BatchedSolver
    b = synth.make_batch(B, N, Jr, Jc, seed=2024)
    P = b["coeffs"].shape[1]
    args = (B, N, N, Jr, Jc, b["coeffs"], P, b["t"], b["diag"], b["y"])
    ref = oracle.best().batched_loglik(*args, nthreads=8)
    alt = oracle.Port().batched_loglik(*args, nthreads=8)
    ll, logdet, quad, status = BatchedSolver().log_likelihood(b["coeffs"], b["t"], b["diag"], b["y"], Jr, Jc,
                                                              return_parts=True)
    assert np.all(status == 0) and np.all(ref[:, 2] == 0)
    c = 0.5 * N * np.log(2 * np.pi)
    for what, got, r, a in (("logdet", logdet, ref[:, 0], alt[:, 0]), ("quad", quad, ref[:, 1], alt[:, 1]),
                            ("loglik", ll, -0.5 * (ref[:, 0] + ref[:, 1]) - c, -0.5 * (alt[:, 0] + alt[:, 1]) - c)):
        worst, msg = scalar_parity(got, r, N, a)
        print("%s %s over %d series: %s" % (name, what, B, msg))
        assert worst <= 1.0, (what, msg)


@pytest.mark.parametrize("Jc", [4, 16])
def test_cfg3_series_with_gradient(Jc):
    """One series of config 3 (N = 100,000) at J = 8 and J = 32: value and all 1+4*Jc derivatives."""
    from celerite_b200 import solver, synth
    N = 100_000
    b = synth.make_batch(1, N, 0, Jc, seed=3100 + Jc)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 0, Jc)
    t, d, y = b["t"][0], b["diag"][0], b["y"][0]
    p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, t, d)
    chk = oracle.best()
    st, rval, rgrad = chk.grad(p, y)
    assert st == 0
    s = solver.CholeskySolver()
    val, g = s.grad_log_likelihood(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, t, y, d)
    const = 0.5 * np.pi * np.log(N)      # solver.cpp:415; the oracle and the drop-in class both carry it
    rv = rval
    e_val, e_grad = abs(val - rv) / abs(rv), rel_grad(g, rgrad)
    print("cfg3 J=%d N=%d: value rel %.3e, gradient worst component %.3e" % (2 * Jc, N, e_val, e_grad))
    assert e_val < RTOL_SCALAR
    assert e_grad < RTOL_GRAD
    # the batched entry point on the same series (the kernel config 3 times)
    from celerite_b200.batched import BatchedSolver
    bv, bg, bst = BatchedSolver().grad_log_likelihood(b["coeffs"], b["t"], b["diag"], b["y"], 0, Jc)
    assert bst[0] == 0
    assert abs((bv[0] - const) - rval) / abs(rval) < RTOL_SCALAR      # the C ABI's batched value carries no constant
    assert rel_grad(bg[0], rgrad) < RTOL_GRAD


def test_cfg4_whole_series():
    """Config 4 at full length: N = 10^7, J = 8.  The reference's own FMA/no-FMA noise on the solve at this
    size is ~1e-10, so vectors are compared norm-wise and the CPU-vs-CPU figure is printed
    and used as the floor of the tolerance."""
    from celerite_b200 import solver, synth
    N, Jc = 10_000_000, 4
    rng = np.random.Generator(np.random.PCG64(42))
    coeffs = synth.make_coeffs(rng, 1, 0, Jc)
    t = np.empty((1, N)); d = np.empty((1, N)); y = np.empty((1, N))
    synth.fill_series(rng, t, d, y)
    z = rng.standard_normal(N)
    t, d, y = t[0], d[0], y[0]
    j, ar, cr, ac, bc, cc, dc = synth.split_row(coeffs[0], 0, Jc)
    p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, t, d)
    chk, port = oracle.best(), oracle.Port()
    f = chk.factor(p)
    zr, lr, qr = chk.solve(f, y)[:, 0], chk.dot_L(f, z)[:, 0], chk.dot_solve(f, y)
    ldr = f.log_det
    if hasattr(chk, "release"):
        chk.release(f)
    fp = port.factor(p)
    zp, lp, qp, ldp = port.solve(fp, y)[:, 0], port.dot_L(fp, z)[:, 0], port.dot_solve(fp, y), fp.log_det
    del fp
    noise = dict(solve=rel_inf(zp, zr), dot_L=rel_inf(lp, lr), quad=abs(qp - qr) / abs(qr), logdet=abs(ldp - ldr) / abs(ldr))
    del zp, lp

    s = solver.CholeskySolver()
    s.compute(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, t, d)
    zg, lg = s.solve(y)[:, 0], s.dot_L(z)[:, 0]
    qg, ldg = s.dot_solve(y), s.log_determinant()
    err = dict(solve=rel_inf(zg, zr), dot_L=rel_inf(lg, lr), quad=abs(qg - qr) / abs(qr), logdet=abs(ldg - ldr) / abs(ldr))
    c = N * np.log(2 * np.pi)
    ll_g, ll_r = -0.5 * (qg + ldg + c), -0.5 * (qr + ldr + c)
    err["loglik"] = abs(ll_g - ll_r) / abs(ll_r)
    noise["loglik"] = abs(-0.5 * (qp + ldp + c) - ll_r) / abs(ll_r)
    for k in ("logdet", "quad", "loglik", "solve", "dot_L"):
        print("cfg4 N=1e7 J=8 %-6s gpu-vs-reference %.3e   reference FMA-vs-noFMA %.3e" % (k, err[k], noise[k]))
    for k, tol in (("logdet", RTOL_SCALAR), ("quad", RTOL_SCALAR), ("loglik", RTOL_SCALAR), ("solve", RTOL_SOLVE),
                   ("dot_L", RTOL_SOLVE)):
        assert err[k] <= max(tol, 3.0 * noise[k]), (k, err[k], noise[k])
