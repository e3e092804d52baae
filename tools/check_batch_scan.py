import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import oracle
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
dev = torch.device("cuda:0")
bs = BatchedSolver(0)
for (B, N, Jr, Jc) in [(1024, 100000, 0, 2), (40, 30000, 1, 1), (100, 8192, 0, 3), (17, 20001, 2, 0)]:
    b = synth.make_batch(B, N, Jr, Jc, seed=B + N)
    tt = lambda a: torch.from_numpy(a).to(dev)
    C, T, D, Y = tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"])
    out = bs.grad_log_likelihood(C, T, D, Y, Jr, Jc); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); bs.grad_log_likelihood(C, T, D, Y, Jr, Jc); e1.record(); torch.cuda.synchronize()
    val, grad, st = out[0].cpu().numpy(), out[1].cpu().numpy(), out[2].cpu().numpy()
    idx = np.unique(np.linspace(0, B - 1, 6).astype(int))
    sub = {k: np.ascontiguousarray(b[k][idx]) for k in ("coeffs", "t", "diag", "y")}
    rv, rg, rs = oracle.best().batched_grad(len(idx), N, N, Jr, Jc, sub["coeffs"], sub["coeffs"].shape[1], sub["t"], sub["diag"], sub["y"], 8)
    rv = rv + 0.5 * np.pi * np.log(N)
    print(B, N, Jr, Jc, "%.2f ms" % e0.elapsed_time(e1), "status", st.sum(), "val rel %.2e" % np.max(np.abs(val[idx] - rv) / np.abs(rv)),
          "grad rel %.2e" % np.max(np.max(np.abs(grad[idx] - rg), axis=1) / np.max(np.abs(rg), axis=1)))
