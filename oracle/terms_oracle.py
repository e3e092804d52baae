"""TEST INFRASTRUCTURE — numpy restatement of the reference's kernel-term -> coefficient maps.

Only tests/ may import this.  Each branch cites the reference lines it follows
(/root/reference/celerite/terms.py).  Pinned against tests/golden/terms_v1.npz, which the
reference itself produced (tests/golden/make_golden_terms.py).

kinds: 0 JitterTerm (327-357), 1 RealTerm (360-391), 2 ComplexTerm without b, 3 ComplexTerm with b
(394-451), 4 SHOTerm with Q >= 1/2, 5 SHOTerm with Q < 1/2 (464-517), 6 Matern32Term (520-566).
A row is [jitter, a_real.., c_real.., a_comp.., b_comp.., c_comp.., d_comp..] with the terms in the
order given, as TermSum.get_all_coefficients concatenates them (terms.py:305-314).
"""
import numpy as np

NPAR = {0: 1, 1: 2, 2: 3, 3: 4, 4: 3, 5: 3, 6: 2, 7: 0}
PRODUCT = 7   # followed in the list by its two factor kinds (TermProduct, terms.py:218-277)


def n_params(kinds):
    return sum(NPAR[k] for k in kinds)


def _factor(k, q, eps):
    """(a, b, c, d, is_complex) of a one-coefficient-set term."""
    if k == 1:
        return np.exp(q[0]), 0.0, np.exp(q[1]), 0.0, False
    if k == 2:
        return np.exp(q[0]), 0.0, np.exp(q[1]), np.exp(q[2]), True
    if k == 3:
        return np.exp(q[0]), np.exp(q[1]), np.exp(q[2]), np.exp(q[3]), True
    if k == 4:
        S0, Q, w0 = np.exp(q[0]), np.exp(q[1]), np.exp(q[2])
        f = np.sqrt(4.0 * Q ** 2 - 1)
        return S0 * w0 * Q, S0 * w0 * Q / f, 0.5 * w0 / Q, 0.5 * w0 / Q * f, True
    if k == 6:
        w0 = np.sqrt(3.0) * np.exp(-q[1])
        S0 = np.exp(2.0 * q[0]) / w0
        return w0 * S0, w0 * w0 * S0 / eps, w0, eps, True
    raise ValueError("not a factor kind")


def coefficients(kinds, eps, params):
    """One parameter vector -> (row, Jr, Jc)."""
    ar, cr, ac, bc, cc, dc = [], [], [], [], [], []
    jitter = 0.0
    o = 0
    t = -1
    kinds = list(kinds)
    while t + 1 < len(kinds):
        t += 1
        k = kinds[t]
        if k == PRODUCT:                             # terms.py:236-277
            k1, k2 = kinds[t + 1], kinds[t + 2]
            q1 = params[o:o + NPAR[k1]]; o += NPAR[k1]
            q2 = params[o:o + NPAR[k2]]; o += NPAR[k2]
            a1, b1, c1, d1, x1 = _factor(k1, q1, eps[t + 1])
            a2, b2, c2, d2, x2 = _factor(k2, q2, eps[t + 2])
            if not x1 and not x2:
                ar.append(a1 * a2); cr.append(c1 + c2)
            elif x1 != x2:
                (aR, cR), (aX, bX, cX, dX) = ((a1, c1), (a2, b2, c2, d2)) if x2 else ((a2, c2), (a1, b1, c1, d1))
                ac.append(aR * aX); bc.append(aR * bX); cc.append(cR + cX); dc.append(dX)
            else:
                ac += [0.5 * (a1 * a2 + b1 * b2), 0.5 * (a1 * a2 - b1 * b2)]
                bc += [0.5 * (b1 * a2 - a1 * b2), 0.5 * (b1 * a2 + a1 * b2)]
                cc += [c1 + c2, c1 + c2]
                dc += [d1 - d2, d1 + d2]
            t += 2
            continue
        q = params[o:o + NPAR[k]]
        o += NPAR[k]
        if k == 0:                                   # terms.py:356-357
            jitter += np.exp(2.0 * q[0])
        elif k == 1:                                 # terms.py:389-391
            ar.append(np.exp(q[0])); cr.append(np.exp(q[1]))
        elif k == 2:                                 # terms.py:431-435
            ac.append(np.exp(q[0])); bc.append(0.0); cc.append(np.exp(q[1])); dc.append(np.exp(q[2]))
        elif k == 3:                                 # terms.py:436-439
            ac.append(np.exp(q[0])); bc.append(np.exp(q[1])); cc.append(np.exp(q[2])); dc.append(np.exp(q[3]))
        elif k == 4:                                 # terms.py:503-517
            S0, Q, w0 = np.exp(q[0]), np.exp(q[1]), np.exp(q[2])
            f = np.sqrt(4.0 * Q ** 2 - 1)
            ac.append(S0 * w0 * Q); bc.append(S0 * w0 * Q / f); cc.append(0.5 * w0 / Q); dc.append(0.5 * w0 / Q * f)
        elif k == 5:                                 # terms.py:490-501
            S0, Q, w0 = np.exp(q[0]), np.exp(q[1]), np.exp(q[2])
            f = np.sqrt(1.0 - 4.0 * Q ** 2)
            ar += [0.5 * S0 * w0 * Q * (1.0 + 1.0 / f), 0.5 * S0 * w0 * Q * (1.0 - 1.0 / f)]
            cr += [0.5 * w0 / Q * (1.0 - f), 0.5 * w0 / Q * (1.0 + f)]
        elif k == 6:                                 # terms.py:562-566
            w0 = np.sqrt(3.0) * np.exp(-q[1])
            S0 = np.exp(2.0 * q[0]) / w0
            ac.append(w0 * S0); bc.append(w0 * w0 * S0 / eps[t]); cc.append(w0); dc.append(eps[t])
        else:
            raise ValueError("unknown term kind")
    row = np.array([jitter] + ar + cr + ac + bc + cc + dc, dtype=np.float64)
    return row, len(ar), len(ac)


def jacobian_fd(kinds, eps, params, h=1e-6):
    """d row / d params by central differences: (n_params, len(row))."""
    params = np.asarray(params, dtype=np.float64)
    out = []
    for i in range(params.size):
        e = np.zeros_like(params)
        e[i] = h
        out.append((coefficients(kinds, eps, params + e)[0] - coefficients(kinds, eps, params - e)[0]) / (2 * h))
    return np.array(out)
