import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
B, N = 65536, 2000
rng = np.random.Generator(np.random.PCG64(42))
c = synth.make_coeffs(rng, B, 0, 1)
t = np.empty((B, N)); d = np.empty((B, N)); y = np.empty((B, N))
synth.fill_series(rng, t, d, y)
dev = torch.device("cuda:0")
tt = lambda a: torch.from_numpy(a).to(dev)
C, T, D, Y = tt(c), tt(t), tt(d), tt(y)
bs = BatchedSolver(0)
ll0, ld0, q0, st0 = bs.log_likelihood(C, T, D, Y, 0, 1, return_parts=True)
torch.cuda.synchronize()
print("first: status sum", int(st0.sum()))
bad = 0
for it in range(int(sys.argv[1]) if len(sys.argv) > 1 else 200):
    ll, ld, q, st = bs.log_likelihood(C, T, D, Y, 0, 1, return_parts=True)
    torch.cuda.synchronize()
    if not (torch.equal(ld, ld0) and torch.equal(q, q0) and torch.equal(st, st0)):
        idx = torch.nonzero((ld != ld0) | (q != q0) | (st != st0)).flatten()
        rel = ((ld - ld0).abs() / ld0.abs())[idx]
        relq = ((q - q0).abs() / q0.abs())[idx]
        if bad < 6:
            print("iter", it, "mismatch in", idx.numel(), "series; first", idx[:8].tolist(), "max rel logdet %.2e quad %.2e" % (float(rel.max()), float(relq.max())),
                  "warps", sorted(set((idx // 32).tolist()))[:8])
        bad += 1
print("bad iterations:", bad)
