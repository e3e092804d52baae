"""TEST INFRASTRUCTURE — numpy restatement of the reference's kernel-term -> coefficient maps.

Only tests/ may import this.  Each branch cites the reference lines it follows
(/root/reference/celerite/terms.py).  Pinned against tests/golden/terms_v1.npz, which the
reference itself produced (tests/golden/make_golden_terms.py).

kinds: 0 JitterTerm (327-357), 1 RealTerm (360-391), 2 ComplexTerm without b, 3 ComplexTerm with b
(394-451), 4 SHOTerm with Q >= 1/2, 5 SHOTerm with Q < 1/2 (464-517), 6 Matern32Term (520-566),
7 TermProduct (218-277) and 8 a sum inside a product, both followed by their two operands (prefix order).
A row is [jitter, a_real.., c_real.., a_comp.., b_comp.., c_comp.., d_comp..] with the terms in the
order given, as TermSum.get_all_coefficients concatenates them (terms.py:305-314).
"""
import numpy as np

NPAR = {0: 1, 1: 2, 2: 3, 3: 4, 4: 3, 5: 3, 6: 2, 7: 0, 8: 0}
PRODUCT = 7   # prefix: followed by its two operands (TermProduct, terms.py:218-277)
SUM = 8       # k1 + k2 as an operand of a product


def n_params(kinds):
    return sum(NPAR[k] for k in kinds)


def _leaf(k, q, eps):
    """a simple term -> (real sets [(a, c)], complex sets [(a, b, c, d)])"""
    if k == 1:                                       # terms.py:389-391
        return [(np.exp(q[0]), np.exp(q[1]))], []
    if k == 2:                                       # terms.py:431-435
        return [], [(np.exp(q[0]), 0.0, np.exp(q[1]), np.exp(q[2]))]
    if k == 3:                                       # terms.py:436-439
        return [], [(np.exp(q[0]), np.exp(q[1]), np.exp(q[2]), np.exp(q[3]))]
    if k == 4:                                       # terms.py:503-517
        S0, Q, w0 = np.exp(q[0]), np.exp(q[1]), np.exp(q[2])
        f = np.sqrt(4.0 * Q ** 2 - 1)
        return [], [(S0 * w0 * Q, S0 * w0 * Q / f, 0.5 * w0 / Q, 0.5 * w0 / Q * f)]
    if k == 5:                                       # terms.py:490-501
        S0, Q, w0 = np.exp(q[0]), np.exp(q[1]), np.exp(q[2])
        f = np.sqrt(1.0 - 4.0 * Q ** 2)
        return [(0.5 * S0 * w0 * Q * (1.0 + 1.0 / f), 0.5 * w0 / Q * (1.0 - f)),
                (0.5 * S0 * w0 * Q * (1.0 - 1.0 / f), 0.5 * w0 / Q * (1.0 + f))], []
    if k == 6:                                       # terms.py:562-566
        w0 = np.sqrt(3.0) * np.exp(-q[1])
        S0 = np.exp(2.0 * q[0]) / w0
        return [], [(w0 * S0, w0 * w0 * S0 / eps, w0, eps)]
    raise ValueError("unknown term kind")


def _expr(kinds, eps, params, pos, o):
    """one operand starting at kinds[pos] with parameters from params[o]: (real, comp, pos, o)"""
    k = kinds[pos]
    if k in (PRODUCT, SUM):
        r1, x1, pos, o = _expr(kinds, eps, params, pos + 1, o)
        r2, x2, pos, o = _expr(kinds, eps, params, pos, o)
        if k == SUM:                                 # terms.py:305-314
            return r1 + r2, x1 + x2, pos, o
        real = [(a1 * a2, c1 + c2) for (a1, c1) in r1 for (a2, c2) in r2]              # terms.py:246-252
        comp = [(aj * ak, aj * bk, cj + ck, dk) for (aj, cj) in r1 for (ak, bk, ck, dk) in x2]   # terms.py:260-267
        comp += [(aj * ak, aj * bk, cj + ck, dk) for (aj, cj) in r2 for (ak, bk, ck, dk) in x1]
        for (aj, bj, cj, dj) in x1:                  # terms.py:269-280
            for (ak, bk, ck, dk) in x2:
                comp.append((0.5 * (aj * ak + bj * bk), 0.5 * (bj * ak - aj * bk), cj + ck, dj - dk))
                comp.append((0.5 * (aj * ak - bj * bk), 0.5 * (bj * ak + aj * bk), cj + ck, dj + dk))
        return real, comp, pos, o
    real, comp = _leaf(k, params[o:o + NPAR[k]], eps[pos])
    return real, comp, pos + 1, o + NPAR[k]


def coefficients(kinds, eps, params):
    """One parameter vector -> (row, Jr, Jc)."""
    kinds = list(kinds)
    real, comp = [], []
    jitter = 0.0
    pos = o = 0
    while pos < len(kinds):
        if kinds[pos] == 0:                          # terms.py:356-357
            jitter += np.exp(2.0 * params[o])
            pos += 1; o += 1
            continue
        r, x, pos, o = _expr(kinds, eps, params, pos, o)
        real += r; comp += x
    row = np.array([jitter] + [r[0] for r in real] + [r[1] for r in real] + [x[0] for x in comp] + [x[1] for x in comp]
                   + [x[2] for x in comp] + [x[3] for x in comp], dtype=np.float64)
    return row, len(real), len(comp)


def jacobian_fd(kinds, eps, params, h=1e-6):
    """d row / d params by central differences: (n_params, len(row))."""
    params = np.asarray(params, dtype=np.float64)
    out = []
    for i in range(params.size):
        e = np.zeros_like(params)
        e[i] = h
        out.append((coefficients(kinds, eps, params + e)[0] - coefficients(kinds, eps, params - e)[0]) / (2 * h))
    return np.array(out)
