"""Generate tests/golden/golden_v1.npz from the UNMODIFIED reference (oracle/_ref/libceleref.so).

Run in the build container (needs /root/reference to build oracle/_ref):
    python tests/golden/make_golden.py
The reference stores no golden vectors of its own, so these fixtures — inputs
plus the reference's outputs for every method on the hot path — are what pins the oracle and
the CUDA path on machines where the reference is not available.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from oracle import Problem  # noqa: E402
from celerite_b200 import synth  # noqa: E402


def cases():
    rng = np.random.default_rng(20261106)
    out = {}

    def series(N, scale=10.0):
        x = np.sort(rng.uniform(0, scale, N))
        diag = rng.uniform(0.1, 0.5, N) ** 2
        y = np.sin(x) + 0.1 * rng.standard_normal(N)
        return x, diag, y

    # cfg1 shape at reduced N: 1 real + 1 complex term (J=3)
    x, d, y = series(512, 40.0)
    out["cfg1_j3"] = (Problem(0.01, [1.3], [0.4], [0.9], [0.05], [0.2], [2.5], x, d), y)
    # cfg2 unit: one SHO term (J=2) on the synthetic TESS-like generator
    b = synth.make_batch(1, 256, 0, 1, seed=7)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 0, 1)
    out["cfg2_sho_j2"] = (Problem(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0]), b["y"][0])
    # cfg5 unit: two SHO terms (J=4)
    b = synth.make_batch(1, 200, 0, 2, seed=8)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 0, 2)
    out["cfg5_sho_j4"] = (Problem(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0]), b["y"][0])
    # the four kernels of cpp/src/test_solvers.cc:79-97 (real, complex, mixed, mixed+general)
    x, d, y = series(128)
    out["real_j2"] = (Problem(0.01, [1.0, 0.6], [0.5, 1.5], [], [], [], [], x, d), y)
    out["complex_j4"] = (Problem(0.01, [], [], [1.0, 0.6], [0.1, 0.2], [0.5, 1.5], [1.0, 4.0], x, d), y)
    out["mixed_j6"] = (Problem(0.01, [1.0, 0.6], [0.5, 1.5], [1.0, 0.6], [0.1, 0.2], [0.5, 1.5], [1.0, 4.0], x, d), y)
    Jg = 3
    U = 0.1 * rng.standard_normal((Jg, 128))
    V = 0.1 * rng.standard_normal((Jg, 128))
    A = 1.0 + np.sum(U * U + V * V, axis=0)
    out["mixed_general_j9"] = (Problem(0.01, [1.0, 0.6], [0.5, 1.5], [1.0, 0.6], [0.1, 0.2], [0.5, 1.5],
                                       [1.0, 4.0], x, d, A=A, U=U, V=V), y)
    # wide kernel: 4 real + 11 complex (J=26), the widest of tests/test_celerite.py:341-370
    x, d, y = series(96)
    out["wide_j26"] = (Problem(0.0, rng.uniform(0.5, 1.5, 4), rng.uniform(0.1, 2, 4),
                               rng.uniform(0.5, 1.5, 11), rng.uniform(0, 0.05, 11),
                               rng.uniform(0.2, 2, 11), rng.uniform(0.5, 8, 11), x, d), y)
    # J=32: 16 SHO terms (cfg3's widest)
    b = synth.make_batch(1, 96, 0, 16, seed=9)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 0, 16)
    out["cfg3_sho_j32"] = (Problem(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0]), b["y"][0])
    # jitter only (J=0): cholesky.h:90-95, 226-231, 334-339, 477-481
    x, d, y = series(16)
    out["jitter_only_j0"] = (Problem(0.3, [], [], [], [], [], [], x, d), y)
    # duplicate time stamp (dt = 0): tests/test_celerite.py:393-404
    x, d, y = series(32)
    x[11] = x[10]
    out["zero_dt_j3"] = (Problem(0.01, [1.3], [0.4], [0.9], [0.05], [0.2], [2.5], x, d), y)
    # tiny N
    x, d, y = series(1)
    out["n1_j3"] = (Problem(0.01, [1.3], [0.4], [0.9], [0.05], [0.2], [2.5], x, d), y)
    x, d, y = series(2)
    out["n2_j3"] = (Problem(0.01, [1.3], [0.4], [0.9], [0.05], [0.2], [2.5], x, d), y)
    # not positive definite: negative amplitude drives D_n < 0 (cholesky.h:176)
    x, d, y = series(64)
    out["npd_j1"] = (Problem(0.0, [-1.0], [0.5], [], [], [], [], x, 1e-3 * np.ones(64)), y)
    # wider than one row per lane (J > 32: cholesky.h:181-204 takes its dynamic branch): 8 real + 16 complex
    # (J=40), the same with three general rows on top of 2 real + 16 complex (J=37), and 32 SHO terms (J=64).
    # Appended last so that the random streams of the cases above stay what they were.
    x, d, y = series(160, 16.0)
    out["wider_j40"] = (Problem(0.01, rng.uniform(0.5, 1.5, 8), rng.uniform(0.1, 2, 8),
                                rng.uniform(0.5, 1.5, 16), rng.uniform(0, 0.05, 16),
                                rng.uniform(0.2, 2, 16), rng.uniform(0.5, 8, 16), x, d), y)
    x, d, y = series(96)
    Jg = 3
    U = 0.1 * rng.standard_normal((Jg, 96))
    V = 0.1 * rng.standard_normal((Jg, 96))
    A = 1.0 + np.sum(U * U + V * V, axis=0)
    out["wider_general_j37"] = (Problem(0.02, rng.uniform(0.5, 1.5, 2), rng.uniform(0.1, 2, 2),
                                        rng.uniform(0.5, 1.5, 16), rng.uniform(0, 0.05, 16),
                                        rng.uniform(0.2, 2, 16), rng.uniform(0.5, 8, 16), x, d, A=A, U=U, V=V), y)
    b = synth.make_batch(1, 64, 0, 32, seed=10)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 0, 32)
    out["wider_sho_j64"] = (Problem(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0]), b["y"][0])
    return out


def main():
    ref = oracle.Reference()
    rng = np.random.default_rng(5)
    blob = {}
    names = []
    for name, (p, y) in cases().items():
        names.append(name)
        k = name + "/"
        blob[k + "jitter"] = np.array(p.jitter)
        for key in ("a_real", "c_real", "a_comp", "b_comp", "c_comp", "d_comp", "x", "diag"):
            blob[k + key] = getattr(p, key)
        if p.Jg:
            blob[k + "A"], blob[k + "U"], blob[k + "V"] = p.A, p.U, p.V
        blob[k + "y"] = y
        f = ref.factor(p)
        blob[k + "status"] = np.array(f.status)
        if f.status != 0:
            st, _, _ = ref.grad(p, y)
            blob[k + "grad_status"] = np.array(st)
            continue
        B = rng.standard_normal((p.N, 3))
        xs = np.sort(rng.uniform(p.x[0] - 1.0, p.x[-1] + 1.0, 40))
        blob[k + "phi"], blob[k + "u"], blob[k + "W"], blob[k + "D"] = f.phi, f.u, f.W, f.D
        blob[k + "log_det"] = np.array(f.log_det)
        blob[k + "dot_solve"] = np.array(ref.dot_solve(f, y))
        blob[k + "rhs"] = B
        blob[k + "solve"] = ref.solve(f, B)
        blob[k + "solve_y"] = ref.solve(f, y)
        blob[k + "dot_L"] = ref.dot_L(f, B)
        blob[k + "dot"] = ref.dot(p, B)
        blob[k + "xs"] = xs
        blob[k + "predict"] = ref.predict(p, f, y, xs)
        st, val, g = ref.grad(p, y)
        blob[k + "grad_status"] = np.array(st)
        blob[k + "grad_val"] = np.array(val)
        blob[k + "grad"] = g
        ref.release(f)
    blob["names"] = np.array(names)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_v1.npz")
    np.savez_compressed(path, **blob)
    print("wrote", path, os.path.getsize(path), "bytes;", len(names), "cases")


if __name__ == "__main__":
    main()
