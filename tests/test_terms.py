"""Kernel-term -> coefficient maps: the numpy restatement against what the
reference's terms.py produced (CPU), and the CUDA maps + transposed Jacobian against both (GPU)."""
import os

import numpy as np
import pytest

from oracle import terms_oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "terms_v1.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def _specs(g):
    for i in range(int(g["nspec"])):
        yield (g["spec%d_kinds" % i], g["spec%d_eps" % i], g["spec%d_params" % i], g["spec%d_rows" % i],
               g["spec%d_jac" % i])


def test_oracle_maps_match_the_reference(gold):
    for kinds, eps, params, rows, jac in _specs(gold):
        for p, row in zip(params, rows):
            got, Jr, Jc = terms_oracle.coefficients(list(kinds), eps, p)
            assert got.shape == row.shape == (1 + 2 * Jr + 4 * Jc,)
            assert np.max(np.abs(got - row) / np.maximum(1e-300, np.abs(row) + 1e-12)) < 1e-13


def test_term_layout_without_a_gpu(gold):
    from celerite_b200.gp import TermSum
    for kinds, eps, params, rows, jac in _specs(gold):
        spec = TermSum(list(kinds))
        assert spec.n_params == params.shape[1] == terms_oracle.n_params(list(kinds))
        assert spec.n_coeffs == rows.shape[1]
        assert len(spec.parameter_names) == spec.n_params
    with pytest.raises(ValueError):
        TermSum([99])
    with pytest.raises(ValueError):                      # products are not implemented for terms with jitter (terms.py:221-224)
        TermSum([("product", "jitter", "real")])
    with pytest.raises(ValueError):                      # a product needs two operands
        TermSum([7, 1])
    nested = TermSum([("product", ("sum", "real", "sho"), "sho_overdamped"), "jitter"])
    assert (nested.n_params, nested.Jr, nested.Jc) == (9, 2, 2)
    assert nested.parameter_names[:3] == ("terms[0]:k1:terms[0]:log_a", "terms[0]:k1:terms[0]:log_c", "terms[0]:k1:terms[1]:log_S0")
    assert nested.parameter_names[-1] == "terms[1]:log_sigma"


@pytest.mark.gpu
def test_cuda_maps_and_backprop(gold):
    import torch
    from celerite_b200.gp import BatchedGP, TermSum
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(4)
    for kinds, eps, params, rows, jac in _specs(gold):
        spec = TermSum(list(kinds), eps={i: e for i, e in enumerate(eps)})
        gp = BatchedGP(spec)
        P = torch.from_numpy(np.ascontiguousarray(params)).to(dev)
        got, status = gp.coefficients(P)
        torch.cuda.synchronize()
        assert int(status.abs().sum()) == 0
        got = got.cpu().numpy()
        assert np.max(np.abs(got - rows) / (np.abs(rows) + 1e-12)) < 1e-13
        # transposed Jacobian: g_params = J g_coeffs with J from central differences of the reference
        gc = rng.standard_normal(rows.shape)
        gpar = gp.backprop(P, torch.from_numpy(gc).to(dev)).cpu().numpy()
        ref = np.einsum("bpc,bc->bp", jac, gc)
        assert np.max(np.abs(gpar - ref)) / np.max(np.abs(ref)) < 1e-7


@pytest.mark.gpu
def test_sho_regime_is_checked():
    import torch
    from celerite_b200.gp import BatchedGP, TermSum
    gp = BatchedGP(TermSum(["sho"]))
    P = torch.tensor([[0.0, np.log(2.0), 0.3], [0.0, np.log(0.2), 0.3]], dtype=torch.float64, device="cuda:0")
    rows, status = gp.coefficients(P)
    assert status.cpu().tolist() == [0, 1]
    assert bool(torch.isnan(rows[1, 1:]).all()) and bool(torch.isfinite(rows[0]).all())


@pytest.mark.gpu
def test_batched_gp_against_the_reference_gp():
    """Walkers of one SHO + real + jitter kernel on one light curve: log-likelihood against the
    reference's GP.log_likelihood, gradient against central differences of it."""
    import sys
    import torch
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref_dir = os.path.join(root, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "celerite")):
        pytest.skip("baseline/_ref (pip install of the reference) is not present")
    sys.path.insert(0, ref_dir)
    try:
        import celerite
        from celerite import terms
    finally:
        sys.path.remove(ref_dir)
    from celerite_b200.gp import BatchedGP, TermSum
    rng = np.random.default_rng(11)
    N, B = 300, 6
    t = np.sort(rng.uniform(0, 30, N))
    yerr = rng.uniform(0.05, 0.2, N)
    y = np.sin(t) + yerr * rng.standard_normal(N)
    params = np.column_stack([rng.uniform(-1, 1, B), np.log(rng.uniform(0.7, 4, B)), rng.uniform(0, 2, B),
                              rng.uniform(-1, 0.5, B), rng.uniform(-2, 0, B), rng.uniform(-4, -2, B)])

    def ref_ll(p):
        k = terms.SHOTerm(log_S0=p[0], log_Q=p[1], log_omega0=p[2]) + terms.RealTerm(log_a=p[3], log_c=p[4]) \
            + terms.JitterTerm(log_sigma=p[5])
        gp = celerite.GP(k)
        gp.compute(t, yerr)
        return gp.log_likelihood(y)

    dev = torch.device("cuda:0")
    bgp = BatchedGP(TermSum(["sho", "real", "jitter"]))
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(np.broadcast_to(a, (B, N)))).to(dev)  # noqa: E731
    P = torch.from_numpy(params).to(dev)
    ll = bgp.log_likelihood(P, tt(t), tt(y), tt(yerr)).cpu().numpy()
    ll2, grad = bgp.grad_log_likelihood(P, tt(t), tt(y), tt(yerr))
    ll2, grad = ll2.cpu().numpy(), grad.cpu().numpy()
    for b in range(B):
        ref = ref_ll(params[b])
        assert abs(ll[b] - ref) / max(abs(ref), N) < 1e-10
        assert abs(ll2[b] - ref) / max(abs(ref), N) < 1e-10
        for i in range(params.shape[1]):
            d = np.zeros(params.shape[1])
            d[i] = 1e-5
            fd = (ref_ll(params[b] + d) - ref_ll(params[b] - d)) / 2e-5
            assert abs(grad[b, i] - fd) < 1e-5 * max(1.0, np.max(np.abs(grad[b])))
