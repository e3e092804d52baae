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
