"""Gradient time on wide batches (many short series) for J beyond the lockstep kernel (development aid)."""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
B, N = 32768, 512
dev = torch.device("cuda:0")
bs = BatchedSolver(0)
for jc in (3, 4, 8, 16):
    b = synth.make_batch(B, N, 0, jc, seed=jc)
    tt = lambda a: torch.from_numpy(a).to(dev)
    C, T, D, Y = tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"])
    def timeit(fn, reps=3):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    tf = timeit(lambda: bs.log_likelihood(C, T, D, Y, 0, jc))
    tg = timeit(lambda: bs.grad_log_likelihood(C, T, D, Y, 0, jc))
    print("wide B=%d N=%d J=%2d forward %.2f ms  grad %.2f ms (%.2e pt-terms/s)" % (B, N, 2 * jc, tf, tg, B * N * 2 * jc / tg * 1e3))
