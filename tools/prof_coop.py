import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
B, N, jc = 1024, 20000, int(sys.argv[1]) if len(sys.argv) > 1 else 4
dev = torch.device("cuda:0")
bs = BatchedSolver(0)
b = synth.make_batch(B, N, 0, jc, seed=jc)
tt = lambda a: torch.from_numpy(a).to(dev)
C, T, D, Y = tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"])
mode = sys.argv[2] if len(sys.argv) > 2 else "ll"
for _ in range(3):
    if mode == "ll": bs.log_likelihood(C, T, D, Y, 0, jc)
    else: bs.grad_log_likelihood(C, T, D, Y, 0, jc)
torch.cuda.synchronize()
print("done")
