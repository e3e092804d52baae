"""Locate the first wrong entry of a chunked sweep (development aid)."""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import oracle
from celerite_b200 import synth, longseries as ls
def run(N, Jr, Jc, nrhs):
    b = synth.make_batch(1, N, Jr, Jc, seed=77)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], Jr, Jc)
    p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0])
    chk = oracle.best(); f = chk.factor(p)
    B = np.random.default_rng(1).standard_normal((N, nrhs))
    whole = ls.LongSeries(0)
    whole.compute(p.jitter, p.a_real, p.c_real, p.a_comp, p.b_comp, p.c_comp, p.d_comp, p.x, p.diag)
    bt = torch.from_numpy(B).cuda()
    for name, got, ref in (("dot_L", whole.dot_L(bt).cpu().numpy(), chk.dot_L(f, B)),):
        err = np.abs(got - ref) / (1 + np.abs(ref))
        bad = np.flatnonzero(~(err.max(axis=1) < 1e-9))
        print(N, Jr, Jc, nrhs, name, "nbad", len(bad), "first", bad[:1], "maxerr", err.max(), "last", bad[-1:])
for N in (60000, 60160, 61440, 61441, 61440 + 128, 30000):
    run(N, 0, 1, 1)
run(61440, 0, 2, 1); run(60160, 0, 2, 1); run(60000, 0, 4, 1); run(60001, 0, 4, 1); run(60001, 1, 1, 2); run(60003, 2, 6, 3)
