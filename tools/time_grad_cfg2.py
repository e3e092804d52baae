"""cfg2 shape (65,536 series x 2,000 points, one SHO term) with the gradient: device time per call and
the largest difference from the oracle on a few series (development aid).
  python tools/time_grad_cfg2.py [B] [N] [jc] [reps]"""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
jc = int(sys.argv[3]) if len(sys.argv) > 3 else 1
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
dev = torch.device("cuda:0")
bs = BatchedSolver(0)
b = synth.make_batch(B, N, 0, jc, seed=7)
tt = lambda a: torch.from_numpy(a).to(dev)
C, T, D, Y = tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"])
out = bs.grad_log_likelihood(C, T, D, Y, 0, jc)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ts = []
for _ in range(reps):
    e0.record(); bs.grad_log_likelihood(C, T, D, Y, 0, jc); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print("grad B=%d N=%d J=%d: min %.3f ms  median %.3f ms" % (B, N, 2 * jc, min(ts), sorted(ts)[len(ts) // 2]))
if os.environ.get("CHECK", "1") == "1":
    import oracle
    chk = oracle.best()
    val, grad = out[0].cpu().numpy(), out[1].cpu().numpy()
    idx = np.arange(0, B, max(1, B // 32))[:32]
    sub = {k: np.ascontiguousarray(b[k][idx]) for k in ("coeffs", "t", "diag", "y")}
    rv, rg, rs = chk.batched_grad(len(idx), N, N, 0, jc, sub["coeffs"], sub["coeffs"].shape[1], sub["t"], sub["diag"], sub["y"], 8)
    has_const = abs((rv[0] - val[idx[0]]) + 0.5 * np.pi * np.log(N)) < abs(rv[0] - val[idx[0]])
    if has_const:
        rv = rv + 0.5 * np.pi * np.log(N)
    worst_v = np.max(np.abs(val[idx] - rv) / np.maximum(1.0, np.abs(rv)))
    worst_g = np.max(np.max(np.abs(grad[idx] - rg), axis=1) / np.maximum(1.0, np.max(np.abs(rg), axis=1)))
    print("vs %s: value rel %.2e  grad rel (by max norm) %.2e" % (type(chk).__name__, worst_v, worst_g))
