"""numpy prototype of the EXACT time-parallel adjoint (grad.cu, scan modes): the adjoint recursion of
tools/proto_adjoint.py is affine in its carries (Pbar, gbar), with the homogeneous part a congruence,

    gbar_out = M gbar_in + g0,     Pbar_out = M Pbar_in M^T - 1/2 ((M gbar_in) g0^T + g0 (M gbar_in)^T) + P0,

where, over the steps n = hi-1 .. lo of a span, M = prod_n Phi_n (I - u_n W_n^T) (lowest step leftmost) and
(g0, P0) are the carries a sweep over the span leaves when it starts from zero.  So spans can be swept
side by side: pass 1 (zero carries in) gives M, g0, P0 per span, a fold over the spans from the last to the
first gives every span its true incoming carries, pass 2 repeats the sweep with them and takes the
gradient sums.  Checked here against the sequential sweep (development aid only).
    python tools/proto_adjoint_scan.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.proto_adjoint import generators  # noqa: E402


def forward(phi, u, v, A, y):
    N, J = phi.shape
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
    return P, W, f, D, x


def sweep(lo, hi, gbar, Pbar, tape, gen, t, partner, want_M=False):
    """adjoint over steps hi-1 .. lo; returns carries out, the seven sums, and M if asked"""
    P, W, f, D, x = tape
    phi, u, v, dx = gen
    J = phi.shape[1]
    gbar = gbar.copy(); Pbar = Pbar.copy()
    sums = np.zeros((7, J))          # R1..R6, and sumA in sums[6, 0]
    M = np.eye(J)
    for n in range(hi - 1, lo - 1, -1):
        PbW = Pbar @ W[n]
        Dbar = W[n] @ PbW + 0.5 * x[n] ** 2 / D[n] ** 2 - 0.5 / D[n]
        xbar = -x[n] / D[n] + gbar @ W[n]
        Wbar = 2.0 * D[n] * PbW + gbar * x[n]
        w = Wbar / D[n]
        Dbar -= (Wbar @ W[n]) / D[n]
        sums[5] += t[n] * w * v[n][partner]
        sums[6, 0] += Dbar
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
        sums[0] += ubar * v[n]
        sums[1] += ubar * v[n][partner]
        sums[2] += phibar * (-dx[n] * phi[n])
        sums[3] += t[n] * ubar * v[n]
        sums[4] += t[n] * ubar * v[n][partner]
        if want_M:
            M = (phi[n][:, None] * (np.eye(J) - np.outer(u[n], W[n]))) @ M
    return gbar, Pbar, sums, M


if __name__ == "__main__":
    rng = np.random.default_rng(5)
    for Jr, Jc, N, K in [(0, 1, 240, 6), (2, 2, 300, 7), (1, 3, 500, 11)]:
        t = np.sort(rng.uniform(0, 30, N)); diag = rng.uniform(0.1, 0.4, N) ** 2; y = rng.standard_normal(N)
        ar, cr = rng.uniform(0.5, 2, Jr), rng.uniform(0.01, 2, Jr)
        ac, bc, cc, dc = rng.uniform(0.5, 2, Jc), rng.uniform(0, 0.02, Jc), rng.uniform(0.01, 2, Jc), rng.uniform(0.5, 6, Jc)
        phi, u, v, A, dx = generators(1e-3, ar, cr, ac, bc, cc, dc, t, diag)
        J = Jr + 2 * Jc
        partner = np.arange(J)
        for q in range(Jc):
            k = Jr + 2 * q
            partner[k], partner[k + 1] = k + 1, k
        tape = forward(phi, u, v, A, y)
        gen = (phi, u, v, dx)
        z, Z = np.zeros(J), np.zeros((J, J))
        g_seq, P_seq, sums_seq, _ = sweep(0, N, z, Z, tape, gen, t, partner)
        edges = np.linspace(0, N, K + 1).astype(int)
        # pass 1: every span from zero carries
        info = [sweep(edges[c], edges[c + 1], z, Z, tape, gen, t, partner, want_M=True) for c in range(K)]
        # fold, last span first
        g_in = [None] * K; P_in = [None] * K
        g, Pb = z, Z
        for c in range(K - 1, -1, -1):
            g_in[c], P_in[c] = g, Pb
            g0, P0, _, M = info[c]
            Mg = M @ g
            Pb = M @ Pb @ M.T - 0.5 * (np.outer(Mg, g0) + np.outer(g0, Mg)) + P0
            g = Mg + g0
        # pass 2: true carries in, sums added up
        sums = np.zeros_like(sums_seq)
        for c in range(K):
            _, _, s, _ = sweep(edges[c], edges[c + 1], g_in[c], P_in[c], tape, gen, t, partner)
            sums += s
        print(Jr, Jc, N, K, "sums: max rel diff to the sequential sweep %.2e" %
              (np.max(np.abs(sums - sums_seq)) / np.max(np.abs(sums_seq))),
              " final carries %.2e" % (np.max(np.abs(g - g_seq)) / max(np.max(np.abs(g_seq)), 1e-300)))
