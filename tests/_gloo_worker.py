"""Worker of tests/test_multiprocess_cpu.py: one rank of a world_size-N gloo group (CPU)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from celerite_b200 import longseries as ls
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    J, nrhs = 5, 2
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ok = True
    for op in (ls.OP_SOLVE_FORWARD, ls.OP_DOT_L, ls.OP_SOLVE_BACKWARD):
        nb = 1 if op == ls.OP_DOT_L else J
        rng = np.random.default_rng(100 * op + rank)          # every rank knows every rank's seed
        mine = torch.from_numpy(rng.standard_normal((nrhs + nb, J)))
        gathered = torch.empty((world * (nrhs + nb), J), dtype=torch.float64)
        dist.all_gather_into_tensor(gathered, mine)
        gathered = gathered.view(world, nrhs + nb, J)
        carry = ls.fold_carries(gathered, rank, op, nrhs, J).numpy()
        g = np.zeros((nrhs, J))                                # direct composition from the seeds
        order = range(0, rank) if op != ls.OP_SOLVE_BACKWARD else range(world - 1, rank, -1)
        for r in order:
            rec = np.random.default_rng(100 * op + r).standard_normal((nrhs + nb, J))
            c = rec[:nrhs]
            if op == ls.OP_DOT_L:
                g = g * rec[nrhs][None, :] + c
            else:
                M = rec[nrhs:].T
                g = (M @ g.T).T + c
        ok = ok and bool(np.allclose(carry, g, rtol=1e-13, atol=1e-13))
    lo, hi = ls.span_bounds(1003, world, rank)
    sizes = torch.tensor([hi - lo])
    dist.all_reduce(sizes)
    ok = ok and int(sizes.item()) == 1003
    dist.destroy_process_group()
    print("RANK %d %s" % (rank, "OK" if ok else "FAIL"))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
