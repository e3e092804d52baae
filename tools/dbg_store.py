import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
B, N, Jr, Jc = 40, 64, 0, 1
b = synth.make_batch(B, N, Jr, Jc, seed=1)
dev = torch.device("cuda:0")
tt = lambda a: torch.from_numpy(a).to(dev)
bs = BatchedSolver(0)
out = bs.compute(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), Jr, Jc)
torch.cuda.synchronize()
print("ok", float(out[3].sum()))
