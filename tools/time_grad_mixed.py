"""Batched gradient with real + complex terms (development aid): python tools/time_grad_mixed.py [B] [N]
   (CEL_GRAD_MIXED_OFF=1: the cooperative kernel)"""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
dev = torch.device("cuda:0"); bs = BatchedSolver(0)
for (Jr, Jc) in [(1, 0), (1, 1), (2, 1), (4, 0)]:
    b = synth.make_batch(B, N, Jr, Jc, seed=9)
    tt = lambda a: torch.from_numpy(a).to(dev)
    C, T, D, Y = tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"])
    bs.grad_log_likelihood(C, T, D, Y, Jr, Jc); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(3):
        e0.record(); bs.grad_log_likelihood(C, T, D, Y, Jr, Jc); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    print("Jr=%d Jc=%d (J=%d) B=%d N=%d: %.3f ms" % (Jr, Jc, Jr + 2 * Jc, B, N, min(ts)))
