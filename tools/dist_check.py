"""Multi-GPU self-check of the long-series path (run under torchrun, one rank per GPU):
solve, dot_L, log-determinant and log-likelihood of DistributedLongSeries against one GPU holding
the whole series.   torchrun --nproc-per-node 2 tools/dist_check.py [N]"""
import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np, torch, torch.distributed as dist
from celerite_b200 import synth, longseries as ls

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
N, Jc = int(sys.argv[1]) if len(sys.argv) > 1 else 400000, 4
b = synth.make_batch(1, N, 0, Jc, seed=5)
j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], 0, Jc)
t, d, y = b["t"][0], b["diag"][0], b["y"][0]
halo = ls.safe_halo(float(np.min(cc)), float(np.mean(np.diff(t))), tol=1e-18, cap=N // (2 * world))
eng = ls.DistributedLongSeries(local)
eng.compute(j, ar, cr, ac, bc, cc, dc, t, d, 2 * halo)     # f forgets like prod(phi): twice the S halo
yl = torch.from_numpy(y[eng.lo:eng.hi]).to(dev)
z = eng.solve(yl); lz = eng.dot_L(yl); ld = eng.log_determinant(dev); ll = eng.log_likelihood(yl, N)
whole = ls.LongSeries(local)
whole.compute(j, ar, cr, ac, bc, cc, dc, t, d)
yw = torch.from_numpy(y).to(dev)
zw, lw = whole.solve(yw)[eng.lo:eng.hi], whole.dot_L(yw)[eng.lo:eng.hi]
err = torch.stack([(z - zw).abs().max() / zw.abs().max(), (lz - lw).abs().max() / lw.abs().max(),
                   (ld - whole.log_determinant(dev)).abs() / N, (ll - whole.log_likelihood(yw)).abs() / N])
dist.all_reduce(err, op=dist.ReduceOp.MAX)
if rank == 0:
    print("world %d N %d: max rel err solve %.2e dot_L %.2e logdet/N %.2e loglik/N %.2e" % ((world, N) + tuple(err.tolist())))
    assert float(err.max()) < 1e-10
    print("OK")
dist.destroy_process_group()
