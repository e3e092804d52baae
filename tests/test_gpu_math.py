"""The kernels' own exp / sincos (celerite_b200/csrc/fastmath.cuh) against numpy's libm, through the
diagnostic entry point cel_debug_math: polynomial cores and the table-driven variants of the fused
batched log-likelihood.  Bars: 2 ulp relative for exp; for sin/cos 2 ulp of max(|sin|, |cos|) <= 1, i.e.
4.5e-16 absolute -- far inside the 1e-10 parity bar, and stated here so that a regression shows up as
a number, not as a drifting log-likelihood."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ULP = 2.0 ** -52


def _run(which, x, two):
    import torch
    from celerite_b200 import _capi
    lib = _capi.lib()
    dev = torch.device("cuda:0")
    dx = torch.from_numpy(x).to(dev)
    o1, o2 = torch.empty_like(dx), torch.empty_like(dx)
    torch.cuda.synchronize(dev)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    rc = lib.cel_debug_math(which, x.size, p(dx), p(o1), p(o2) if two else None, None)
    assert rc == _capi.OK
    torch.cuda.synchronize(dev)
    return o1.cpu().numpy(), o2.cpu().numpy()


@pytest.mark.parametrize("which", [0, 1])
def test_exp(which):
    rng = np.random.default_rng(1)
    x = np.concatenate([-rng.uniform(0, 30, 400000), -10.0 ** rng.uniform(-12, 2.8, 200000), [0.0, -1e-300, -700.0]])
    got, _ = _run(which, x, False)
    ref = np.exp(x)
    err = np.abs(got - ref) / ref
    print("exp variant %d: max rel err %.2f ulp" % (which, err.max() / ULP))
    assert err.max() < 2.0 * ULP


@pytest.mark.parametrize("which", [0, 1])
def test_sincos(which):
    rng = np.random.default_rng(2)
    x = np.concatenate([rng.uniform(-400, 400, 400000), 10.0 ** rng.uniform(-8, 8.9, 200000) * rng.choice([-1, 1], 200000),
                        [0.0, np.pi / 64, np.pi / 2, 1e-300]])
    s, c = _run(which, x, True)
    es, ec = np.abs(s - np.sin(x)), np.abs(c - np.cos(x))
    print("sincos variant %d: max abs err sin %.2f ulp(1), cos %.2f ulp(1)" % (which, es.max() / ULP, ec.max() / ULP))
    assert max(es.max(), ec.max()) < 2.0 * ULP
    # the pair stays on the unit circle
    assert np.max(np.abs(s * s + c * c - 1.0)) < 4.0 * ULP


def test_reciprocal_of_the_hot_loops():
    rng = np.random.default_rng(3)
    x = np.concatenate([10.0 ** rng.uniform(-150, 150, 300000), rng.uniform(0.5, 2.0, 300000)])
    got, _ = _run(2, x, False)
    err = np.abs(got * x - 1.0)
    print("fast_rcp_cubic: max |x r - 1| = %.2f ulp" % (err.max() / ULP))
    assert err.max() < 2.0 * ULP
