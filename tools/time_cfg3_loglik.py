import sys, os, time
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
bs = BatchedSolver(0)
dev = torch.device("cuda:0")
for Jc in (1, 2, 4, 8, 16):
    B, N = 1024, 100000
    b = synth.make_batch(B, N, 0, Jc, seed=3)
    tt = lambda a: torch.from_numpy(a).to(dev)
    c, t, d, y = tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"])
    for _ in range(2): bs.log_likelihood(c, t, d, y, 0, Jc)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): bs.log_likelihood(c, t, d, y, 0, Jc)
    e1.record(); torch.cuda.synchronize()
    print("loglik B=1024 N=1e5 J=%d: %.2f ms" % (2 * Jc, e0.elapsed_time(e1) / 3))
