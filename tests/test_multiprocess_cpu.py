"""world_size-2 gloo test (CPU) of the multi-GPU host logic of the long-series path: span
partitioning, the all-gather of span summaries and the folding of the gathered affine maps into
each rank's incoming carry (celerite_b200/longseries.py).  The CUDA sweeps themselves are
replaced by random affine maps here; tests/test_gpu_longseries.py checks them on a GPU."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_span_bounds_cover_everything():
    from celerite_b200 import longseries as ls
    for N in (1, 7, 1000, 10**7 + 3):
        for world in (1, 2, 3, 8):
            b = [ls.span_bounds(N, world, r) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == N
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            assert max(h - l for l, h in b) - min(h - l for l, h in b) <= 1


def test_c_abi_span_bounds_agree_with_the_python_statement():
    """cel_mgpu_bounds (the split cel_compute_mgpu uses) is plain integer arithmetic: callable without a GPU."""
    import ctypes as C
    from celerite_b200 import _capi, longseries as ls
    lib = _capi.lib()
    for N in (8, 1000, 10**7 + 3, 2**33 + 5):
        for world in (1, 2, 3, 8):
            for r in range(world):
                lo, hi = C.c_int64(), C.c_int64()
                assert lib.cel_mgpu_bounds(N, world, r, C.byref(lo), C.byref(hi)) == _capi.OK
                assert (lo.value, hi.value) == ls.span_bounds(N, world, r)
    assert lib.cel_mgpu_bounds(10, 2, 2, C.byref(lo), C.byref(hi)) == _capi.INVALID_ARGUMENT


def test_fold_carries_gloo_world2():
    import subprocess
    import sys
    port = _free_port()
    here = os.path.dirname(os.path.abspath(__file__))
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(here, "_gloo_worker.py")], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=280)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "RANK 0 OK" in outs[0] and "RANK 1 OK" in outs[1]


def test_safe_halo():
    from celerite_b200 import longseries as ls
    assert ls.safe_halo(0.5, 0.02) >= 2000
    assert ls.safe_halo(1e-9, 0.02, cap=100) == 100
