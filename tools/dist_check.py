"""Multi-GPU self-check of the long-series path (run under torchrun, one rank per GPU):
solve, dot_L and log-likelihood of one series split over the ranks -- collectives inside the C ABI
(cel_comm_init + cel_*_mgpu) and, for comparison, through torch.distributed -- against one GPU holding
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
whole = ls.LongSeries(local)
whole.compute(j, ar, cr, ac, bc, cc, dc, t, d)
yw = torch.from_numpy(y).to(dev)
rhs = torch.stack([yw, torch.from_numpy(np.random.default_rng(3).standard_normal(N)).to(dev)], dim=1)   # two columns
zw_all, lw_all, llw = whole.solve(rhs), whole.dot_L(rhs), float(whole.log_likelihood(yw))
worst = 0.0
for use_c_abi in (True, False):
    eng = ls.DistributedLongSeries(local, use_c_abi=use_c_abi)
    assert eng.c_abi == use_c_abi
    eng.compute(j, ar, cr, ac, bc, cc, dc, t, d, 2 * halo)     # f forgets like prod(phi): twice the S halo
    rl = rhs[eng.lo:eng.hi].contiguous()
    z, lz = eng.solve(rl), eng.dot_L(rl)
    ll = float(eng.log_likelihood(rl[:, 0].contiguous(), N))
    zw, lw = zw_all[eng.lo:eng.hi], lw_all[eng.lo:eng.hi]
    err = torch.stack([(z - zw).abs().max() / zw.abs().max(), (lz - lw).abs().max() / lw.abs().max(),
                       torch.tensor(abs(ll - llw) / abs(llw), device=dev, dtype=torch.float64)])
    dist.all_reduce(err, op=dist.ReduceOp.MAX)
    worst = max(worst, float(err.max()))
    if rank == 0:
        print("world %d N %d [%s]: max rel err solve %.2e dot_L %.2e loglik %.2e"
              % ((world, N, eng.collective_backend) + tuple(err.tolist())))
    del eng
if rank == 0:
    assert worst < 1e-10
    print("OK")
dist.destroy_process_group()
