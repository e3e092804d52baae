import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
dev = torch.device("cuda:0"); bs = BatchedSolver(0)
for (Jr, Jc) in [(0, 1), (1, 1), (0, 2)]:
    B, N = 65536, 1000
    b = synth.make_batch(B, N, Jr, Jc, seed=3)
    ln = np.random.default_rng(0).integers(N // 2, N + 1, B).astype(np.int32)
    tt = lambda a: torch.from_numpy(a).to(dev)
    C, T, D, Y, L = tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"]), tt(ln)
    bs.log_likelihood(C, T, D, Y, Jr, Jc, lengths=L); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(5):
        e0.record(); bs.log_likelihood(C, T, D, Y, Jr, Jc, lengths=L); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    print("ragged Jr=%d Jc=%d: %.3f ms" % (Jr, Jc, min(ts)))
