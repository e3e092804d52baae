"""First-solve (chunk maps built) vs repeated-solve time on a long series (development aid)."""
import sys, os, time
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth, longseries as ls
N, Jc = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000, 4
b = synth.make_batch(1, N, 0, Jc, seed=3)
j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 0, Jc)
L = ls.LongSeries(0)
rhs = torch.from_numpy(b["y"][0]).cuda()
for rep in range(2):
    t0 = time.perf_counter()
    L.compute(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0])
    torch.cuda.synchronize()
    print("compute (host arrays in, factor resident): %.1f ms wall" % ((time.perf_counter() - t0) * 1e3))
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    ev[0].record(); z = L.solve(rhs); ev[1].record(); z = L.solve(rhs); ev[2].record(); w = L.dot_L(rhs); ev[3].record()
    torch.cuda.synchronize()
    print("rep %d: first solve %.2f ms, second solve %.2f ms, first dot_L %.2f ms" % (rep, ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])))
from celerite_b200 import solver
S = solver.CholeskySolver()
E0, E2 = np.empty(0), np.empty((0, 0))
for rep in range(2):
    t0 = time.perf_counter(); S.compute(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, b["t"][0], b["diag"][0]); t1 = time.perf_counter()
    q = S.dot_solve(b["y"][0]); t2 = time.perf_counter()
    print("host API: compute %.1f ms, dot_solve %.1f ms (log-likelihood of one N=%d series: %.1f ms)" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, N, (t2 - t0) * 1e3))
for rep in range(2):
    t0 = time.perf_counter()
    v, g = S.grad_log_likelihood(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, b["t"][0], b["y"][0], b["diag"][0])
    print("host API: grad_log_likelihood of one N=%d series: %.1f ms (launches so far %d)" % (N, (time.perf_counter() - t0) * 1e3, S.launch_count()))
