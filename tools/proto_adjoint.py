"""numpy prototype of the reverse-mode gradient sweep implemented by celerite_b200/csrc/grad.cu.

Development aid only (not used by the product): it states the adjoint recursion in exactly the
row-wise form the kernel uses and checks it against the oracle's forward-mode gradient.
    python tools/proto_adjoint.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def generators(jitter, ar, cr, ac, bc, cc, dc, t, diag):
    N, Jr, Jc = len(t), len(ar), len(ac)
    J = Jr + 2 * Jc
    phi = np.zeros((N, J)); u = np.zeros((N, J)); v = np.zeros((N, J))
    dx = np.diff(t, prepend=t[0])
    for j in range(Jr):
        phi[:, j] = np.exp(-cr[j] * dx); u[:, j] = ar[j]; v[:, j] = 1.0
    for q in range(Jc):
        k = Jr + 2 * q
        cd, sd = np.cos(dc[q] * t), np.sin(dc[q] * t)
        e = np.exp(-cc[q] * dx)
        phi[:, k] = e; phi[:, k + 1] = e
        u[:, k] = ac[q] * cd + bc[q] * sd; u[:, k + 1] = ac[q] * sd - bc[q] * cd
        v[:, k] = cd; v[:, k + 1] = sd
    A = diag + ar.sum() + ac.sum() + jitter
    return phi, u, v, A, dx


def grad(jitter, ar, cr, ac, bc, cc, dc, t, y, diag):
    ar, cr, ac, bc, cc, dc = (np.asarray(a, float) for a in (ar, cr, ac, bc, cc, dc))
    N, Jr, Jc = len(t), len(ar), len(ac)
    J = Jr + 2 * Jc
    phi, u, v, A, dx = generators(jitter, ar, cr, ac, bc, cc, dc, t, diag)
    # ---- forward, keeping the tape {P_n, W_n, f_n, D_n, x_n} ----
    P = np.zeros((N, J, J)); W = np.zeros((N, J)); f = np.zeros((N, J)); D = np.zeros(N); x = np.zeros(N)
    S = np.zeros((J, J))
    D[0] = A[0]; W[0] = v[0] / D[0]; x[0] = y[0]
    for n in range(1, N):
        P[n] = S + D[n - 1] * np.outer(W[n - 1], W[n - 1])
        S = np.outer(phi[n], phi[n]) * P[n]
        su = S @ u[n]
        D[n] = A[n] - u[n] @ su
        W[n] = (v[n] - su) / D[n]
        f[n] = phi[n] * (f[n - 1] + W[n - 1] * x[n - 1])
        x[n] = y[n] - u[n] @ f[n]
    val = -0.5 * (np.sum(x * x / D) + np.sum(np.log(D)))
    # ---- adjoint, n = N-1 .. 0 ----
    gbar = np.zeros(J); Pbar = np.zeros((J, J))      # carries: adjoints of g_n and P_{n+1}
    sumA = 0.0
    R1 = np.zeros(J); R2 = np.zeros(J); R3 = np.zeros(J); R4 = np.zeros(J); R5 = np.zeros(J); R6 = np.zeros(J)
    partner = np.arange(J)
    for q in range(Jc):
        k = Jr + 2 * q
        partner[k], partner[k + 1] = k + 1, k
    for n in range(N - 1, -1, -1):
        PbW = Pbar @ W[n]
        Dbar = W[n] @ PbW + 0.5 * x[n] ** 2 / D[n] ** 2 - 0.5 / D[n]
        xbar = -x[n] / D[n] + gbar @ W[n]
        Wbar = 2.0 * D[n] * PbW + gbar * x[n]
        w = Wbar / D[n]
        Dbar -= (Wbar @ W[n]) / D[n]
        vbar = w
        R6 += t[n] * vbar * v[n][partner]
        sumA += Dbar
        if n == 0:
            break
        fbar = gbar - xbar * u[n]
        ubar = -xbar * f[n]
        gprev = f[n - 1] + W[n - 1] * x[n - 1]
        phibar = fbar * gprev
        gbar = phi[n] * fbar
        Sn = np.outer(phi[n], phi[n]) * P[n]
        ubar -= Sn @ w + 2.0 * Dbar * (Sn @ u[n])
        Sbar = Pbar - 0.5 * (np.outer(w, u[n]) + np.outer(u[n], w)) - Dbar * np.outer(u[n], u[n])
        phibar += 2.0 * (Sbar * P[n]) @ phi[n]
        Pbar = np.outer(phi[n], phi[n]) * Sbar
        R1 += ubar * v[n]
        R2 += ubar * v[n][partner]
        R3 += phibar * (-dx[n] * phi[n])
        R4 += t[n] * ubar * v[n]
        R5 += t[n] * ubar * v[n][partner]
    g = np.zeros(1 + 2 * Jr + 4 * Jc)
    g[0] = sumA if jitter > np.finfo(float).eps else 0.0
    for j in range(Jr):
        g[1 + j] = R1[j] + sumA
        g[1 + Jr + j] = R3[j]
    o = 1 + 2 * Jr
    for q in range(Jc):
        k = Jr + 2 * q
        g[o + q] = R1[k] + R1[k + 1] + sumA
        g[o + Jc + q] = R2[k] - R2[k + 1]
        g[o + 2 * Jc + q] = R3[k] + R3[k + 1]
        g[o + 3 * Jc + q] = (-ac[q] * R5[k] + bc[q] * R4[k] + ac[q] * R5[k + 1] + bc[q] * R4[k + 1]
                             - R6[k] + R6[k + 1])
    return val, g


if __name__ == "__main__":
    import oracle
    rng = np.random.default_rng(2)
    port = oracle.Port()
    for Jr, Jc, N in [(2, 2, 300), (0, 1, 200), (1, 0, 50), (3, 4, 120)]:
        t = np.sort(rng.uniform(0, 20, N)); diag = rng.uniform(0.1, 0.4, N) ** 2; y = rng.standard_normal(N)
        ar, cr = rng.uniform(0.5, 2, Jr), rng.uniform(0.1, 2, Jr)
        ac, bc, cc, dc = rng.uniform(0.5, 2, Jc), rng.uniform(0, 0.02, Jc), rng.uniform(0.1, 2, Jc), rng.uniform(0.5, 6, Jc)
        p = oracle.Problem(1e-3, ar, cr, ac, bc, cc, dc, t, diag)
        st, vref, gref = port.grad(p, y)
        val, g = grad(1e-3, ar, cr, ac, bc, cc, dc, t, y, diag)
        vref_noconst = vref + 0.5 * np.pi * np.log(N)
        print(Jr, Jc, N, "val rel", abs(val - vref_noconst) / abs(vref_noconst),
              "grad max rel", np.max(np.abs(g - gref) / np.maximum(np.abs(gref), 1e-12 * np.abs(gref).max())))
