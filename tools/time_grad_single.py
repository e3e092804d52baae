import sys, os, time
sys.path.insert(0, os.getcwd())
import numpy as np
import oracle
from celerite_b200 import solver, synth
E0, E2 = np.empty(0), np.empty((0, 0))
for (N, Jr, Jc) in [(4096, 1, 1), (2048, 0, 1), (3000, 2, 3), (10000, 0, 4), (2047, 1, 1)]:
    b = synth.make_batch(1, N, Jr, Jc, seed=N)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], Jr, Jc)
    t, d, y = b["t"][0], b["diag"][0], b["y"][0]
    s = solver.CholeskySolver()
    val, g = s.grad_log_likelihood(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, t, y, d)
    n0 = s.launch_count(); ts = []
    for _ in range(20):
        t0 = time.perf_counter(); val, g = s.grad_log_likelihood(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, t, y, d); ts.append(time.perf_counter() - t0)
    nl = (s.launch_count() - n0) // 20
    p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, t, d)
    st, vref, gref = oracle.best().grad(p, y)
    print("N=%d J=%d: %.3f ms, %d launches; val rel %.1e grad rel %.1e" % (N, Jr + 2 * Jc, 1e3 * np.median(ts), nl,
          abs(val - vref) / abs(vref), np.max(np.abs(g - gref) / np.maximum(np.abs(gref), 1e-12 * np.abs(gref).max()))))
