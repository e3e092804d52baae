import os, sys, numpy as np
sys.path.insert(0, os.getcwd())
import oracle
from celerite_b200 import synth
from celerite_b200.batched import BatchedSolver
B, N, Jr, Jc = 16384, 1000, 0, 2
b = synth.make_batch(B, N, Jr, Jc, seed=5151)
P = b["coeffs"].shape[1]
args = (B, N, N, Jr, Jc, b["coeffs"], P, b["t"], b["diag"], b["y"])
ref = oracle.best().batched_loglik(*args, nthreads=16)
alt = oracle.Port().batched_loglik(*args, nthreads=16)
ll, ld, q, st = BatchedSolver().log_likelihood(b["coeffs"], b["t"], b["diag"], b["y"], Jr, Jc, return_parts=True)
for name, got, r, a in (("logdet", ld, ref[:, 0], alt[:, 0]), ("quad", q, ref[:, 1], alt[:, 1])):
    e = np.abs(got - r) / np.abs(r); n = np.abs(a - r) / np.abs(r)
    i = int(np.argmax(e))
    print("%s %s: gpu-vs-ref median %.2e p99 %.2e max %.2e (ref %.4g, abs err %.2e; cpu-vs-cpu there %.2e) | cpu-vs-cpu median %.2e p99 %.2e max %.2e"
          % (os.environ.get("TAG", ""), name, np.median(e), np.percentile(e, 99), e.max(), r[i], abs(got[i] - r[i]), n[i], np.median(n), np.percentile(n, 99), n.max()))
    ea = np.abs(got - r); na = np.abs(a - r)
    print("     abs: gpu median %.2e p99 %.2e max %.2e | cpu-vs-cpu median %.2e p99 %.2e max %.2e" % (np.median(ea), np.percentile(ea, 99), ea.max(), np.median(na), np.percentile(na, 99), na.max()))
