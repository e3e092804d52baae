"""Solve with several right-hand sides on a long series (development aid)."""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth, longseries as ls
N, Jc = 2_000_000, 4
b = synth.make_batch(1, N, 0, Jc, seed=3)
j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 0, Jc)
L = ls.LongSeries(0)
L.compute(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0])
for nrhs in (1, 2, 4, 8):
    B = torch.randn(N, nrhs, dtype=torch.float64, device="cuda")
    L.solve(B); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): L.solve(B)
    e1.record(); torch.cuda.synchronize()
    print("N=%d J=8 solve nrhs=%d: %.2f ms" % (N, nrhs, e0.elapsed_time(e1) / 5))
