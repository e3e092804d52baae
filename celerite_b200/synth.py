"""Synthetic irregularly-sampled series of the shapes BASELINE.json names.

Host-side numpy only; shared by bench.py and the tests so that the GPU path, the oracle and
the CPU baseline all see byte-identical inputs.  Everything is fp64.

    t     : t0 = 0, dt_n = h*(0.1 + 1.8*U(0,1)), h = 0.02 d, 1 % chance of a 50*h gap
    diag  : sigma_n^2, sigma_n ~ U(0.1, 0.5)          (range of tests/test_celerite.py:51,94)
    y     : sin(t_n) + 0.1*N(0,1)                      (sin t as in cpp/src/benchmark.cc:63)
    jitter: 0.01                                       (cpp/src/benchmark.cc:40)
    SHO   : S0~U(0.5,2), Q~U(0.7,5) (> 1/2 => one complex term), w0~LogU(0.5,20),
            coefficients by celerite/terms.py:503-517
    real  : a~U(0.5,2), c~LogU(0.05,5)
"""
import numpy as np

H_CADENCE = 0.02
JITTER = 0.01


def coeff_count(Jr, Jc):
    """Length of one coefficient row [jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp]."""
    return 1 + 2 * Jr + 4 * Jc


def sho_coefficients(S0, Q, w0):
    """(a, b, c, d) of the single complex term of an under-damped SHO (terms.py:503-517)."""
    f = np.sqrt(4.0 * Q ** 2 - 1.0)
    return S0 * w0 * Q, S0 * w0 * Q / f, 0.5 * w0 / Q, 0.5 * w0 / Q * f


def make_coeffs(rng, B, Jr, Jc, jitter=JITTER):
    """(B, P) rows of coefficients: Jr real terms then Jc SHO terms, all positive definite."""
    P = coeff_count(Jr, Jc)
    out = np.empty((B, P))
    out[:, 0] = jitter
    o = 1
    out[:, o:o + Jr] = rng.uniform(0.5, 2.0, (B, Jr)); o += Jr
    out[:, o:o + Jr] = np.exp(rng.uniform(np.log(0.05), np.log(5.0), (B, Jr))); o += Jr
    S0 = rng.uniform(0.5, 2.0, (B, Jc))
    Q = rng.uniform(0.7, 5.0, (B, Jc))
    w0 = np.exp(rng.uniform(np.log(0.5), np.log(20.0), (B, Jc)))
    a, b, c, d = sho_coefficients(S0, Q, w0)
    out[:, o:o + Jc] = a; o += Jc
    out[:, o:o + Jc] = b; o += Jc
    out[:, o:o + Jc] = c; o += Jc
    out[:, o:o + Jc] = d
    return out


def fill_series(rng, t, diag, y):
    """Fill pre-allocated (B, N) float64 arrays in place (they may be pinned torch buffers)."""
    B, N = t.shape
    step = max(1, (1 << 22) // max(N, 1))
    for b0 in range(0, B, step):
        b1 = min(B, b0 + step)
        n = b1 - b0
        dt = H_CADENCE * (0.1 + 1.8 * rng.random((n, N)))
        gap = rng.random((n, N)) < 0.01
        dt[gap] = 50.0 * H_CADENCE
        dt[:, 0] = 0.0
        np.cumsum(dt, axis=1, out=t[b0:b1])
        sig = rng.uniform(0.1, 0.5, (n, N))
        np.multiply(sig, sig, out=diag[b0:b1])
        np.sin(t[b0:b1], out=y[b0:b1])
        y[b0:b1] += 0.1 * rng.standard_normal((n, N))


def make_batch(B, N, Jr, Jc, seed=42, jitter=JITTER):
    """Returns dict(t, diag, y: (B,N); coeffs: (B,P)) for B independent series."""
    rng = np.random.Generator(np.random.PCG64(seed))
    coeffs = make_coeffs(rng, B, Jr, Jc, jitter)
    t = np.empty((B, N)); diag = np.empty((B, N)); y = np.empty((B, N))
    fill_series(rng, t, diag, y)
    return dict(t=t, diag=diag, y=y, coeffs=coeffs, B=B, N=N, Jr=Jr, Jc=Jc)


def split_row(row, Jr, Jc):
    """coefficient row -> (jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp)."""
    o = 1
    parts = []
    for n in (Jr, Jr, Jc, Jc, Jc, Jc):
        parts.append(np.array(row[o:o + n], dtype=np.float64))
        o += n
    return (float(row[0]),) + tuple(parts)
