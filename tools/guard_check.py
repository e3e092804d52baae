"""Small instances of the gradient / term-map / ragged paths under CEL_DEBUG_GUARD: every internal device buffer
is then allocated at exactly the size asked for plus a guard band, and cel_debug_guards_intact says whether a
kernel wrote behind one (tests/test_gpu_batched.py::test_no_kernel_writes_behind_its_scratch runs this):
   CEL_DEBUG_GUARD=1 python tools/guard_check.py"""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from celerite_b200 import synth, solver
from celerite_b200.batched import BatchedSolver
from celerite_b200.gp import BatchedGP, TermSum
dev = torch.device("cuda:0")
bs = BatchedSolver(0)
tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def intact(h, what):
    rc = bs._lib.cel_debug_guards_intact(h)
    assert rc == 1, "guard band touched (or CEL_DEBUG_GUARD not set): %s -> %d" % (what, rc)


for (B, N, Jr, Jc) in [(70, 404, 0, 1), (5, 2002, 0, 1), (33, 18, 0, 1), (6, 8192, 0, 3), (3, 4200, 1, 1), (9, 300, 0, 2),
                       (2100, 100, 1, 1), (2090, 62, 2, 1), (2500, 50, 0, 2)]:
    b = synth.make_batch(B, N, Jr, Jc, seed=B + N)
    out = bs.grad_log_likelihood(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"]), Jr, Jc)
    torch.cuda.synchronize()
    intact(bs._h, ("batched grad", B, N, Jr, Jc))
    v, g, s = bs.grad_log_likelihood(b["coeffs"], b["t"], b["diag"], b["y"], Jr, Jc)     # host arrays
    intact(bs._h, ("batched grad, host arrays", B, N, Jr, Jc))
    assert np.allclose(v, out[0].cpu().numpy(), rtol=1e-12) and np.allclose(g, out[1].cpu().numpy(), rtol=1e-9, atol=1e-9)
    ll = bs.log_likelihood(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), tt(b["y"]), Jr, Jc)
    intact(bs._h, ("batched log-likelihood", B, N, Jr, Jc))
    print("batched grad", B, N, Jr, Jc, float(out[0].sum()))
E0, E2 = np.empty(0), np.empty((0, 0))
for (N, Jc, env) in [(20000, 2, None), (17000, 1, "CEL_GRAD_SCAN_SERIAL_FORWARD")]:
    if env:
        os.environ[env] = "1"
    b = synth.make_batch(1, N, 0, Jc, seed=N)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 0, Jc)
    s = solver.CholeskySolver()
    val, g = s.grad_log_likelihood(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, b["t"][0], b["y"][0], b["diag"][0])
    intact(s._h, ("long series grad", N, Jc, env))
    s.compute(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, b["t"][0], b["diag"][0])
    z = s.solve(b["y"][0]); w = s.dot_L(b["y"][0])
    intact(s._h, ("long series compute / solve / dot_L", N, Jc))
    print("long series grad", N, Jc, val)
gp = BatchedGP(TermSum([("product", ("sum", "real", "sho"), "sho_overdamped"), "jitter"]))
P = torch.tensor(np.random.default_rng(0).uniform(-1, 0.5, (5, 9)))
P[:, 6] = np.log(0.3); P[:, 3] = np.log(2.0)     # log_Q of the over-damped and of the under-damped SHO
rows, st = gp.coefficients(P.to(dev))
gpar = gp.backprop(P.to(dev), torch.ones_like(rows))
torch.cuda.synchronize()
intact(gp._h, "term maps")
assert int(st.sum()) == 0 and bool(torch.isfinite(gpar).all())
print("terms", rows.shape, float(gpar.sum()))
b = synth.make_batch(7, 200, 1, 2, seed=1)
ln = torch.tensor([200, 1, 2, 50, 199, 3, 100], dtype=torch.int32, device=dev)
o = bs.compute(tt(b["coeffs"]), tt(b["t"]), tt(b["diag"]), 1, 2, lengths=ln)
torch.cuda.synchronize()
intact(bs._h, "ragged compute")
print("ragged compute", float(o[4].sum()))
print("guards intact")
