"""Gradient of ONE long series through the host API (development aid): python tools/time_grad_long.py [N] [Jc]"""
import sys, os, time
sys.path.insert(0, os.getcwd())
import numpy as np
from celerite_b200 import solver, synth
N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
Jc = int(sys.argv[2]) if len(sys.argv) > 2 else 4
b = synth.make_batch(1, N, 0, Jc, seed=42)
j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 0, Jc)
if len(sys.argv) > 3:
    cc = cc * float(sys.argv[3])      # slower decay: longer halos
E0, E2 = np.empty(0), np.empty((0, 0))
s = solver.CholeskySolver()
t, d, y = b["t"][0], b["diag"][0], b["y"][0]
print("c_min", cc.min(), "mean dt", (t[-1] - t[0]) / (N - 1))
for rep in range(3):
    n0 = s.launch_count(); t0 = time.perf_counter()
    val, g = s.grad_log_likelihood(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, t, y, d)
    dt = time.perf_counter() - t0
    print("N=%d J=%d: %.1f ms, %d launches, val %.6f" % (N, 2 * Jc, dt * 1e3, s.launch_count() - n0, val))
print("grad[:3]", g[:3])
