"""Batched fused log-likelihood over (Jr, Jc) combinations (development aid):
   python tools/time_loglik_batched.py [B] [N]     (CEL_LOCK_V1=1 for the first-generation loop)"""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import oracle
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
dev = torch.device("cuda:0")
bs = BatchedSolver(0)
for (Jr, Jc) in [(1, 0), (0, 1), (1, 1), (2, 1), (0, 2), (1, 2), (0, 3), (2, 2)]:
    b = synth.make_batch(B, N, Jr, Jc, seed=Jr + 10 * Jc)
    tt = lambda a: torch.from_numpy(a).to(dev)
    C, T, D, Y = tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"])
    out = bs.log_likelihood(C, T, D, Y, Jr, Jc, return_parts=True); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(5):
        e0.record(); bs.log_likelihood(C, T, D, Y, Jr, Jc); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    n = 256
    ref = oracle.best().batched_loglik(n, N, N, Jr, Jc, b["coeffs"][:n].copy(), b["coeffs"].shape[1], b["t"][:n].copy(), b["diag"][:n].copy(), b["y"][:n].copy(), 8)
    ld, q = out[1].cpu().numpy()[:n], out[2].cpu().numpy()[:n]
    print("Jr=%d Jc=%d (J=%d): %.3f ms  (%.0f GB/s)  logdet rel %.1e quad rel %.1e" % (Jr, Jc, Jr + 2 * Jc, min(ts), B * N * 24 / min(ts) / 1e6,
          np.max(np.abs(ld - ref[:, 0]) / np.maximum(1, np.abs(ref[:, 0]))), np.max(np.abs(q - ref[:, 1]) / np.abs(ref[:, 1]))))
