"""The allocation-free gradient entry point, cel_grad_log_likelihood_static_scratch.

Some users of this library run inside a host program that owns all device memory through its
own pool and forbids cudaMalloc/cudaFree once it is up (both stall the whole device), or work
under a fixed memory budget.  For them the library offers one gradient call whose scratch is a
fixed 32 MiB array inside the library image, so the call itself allocates nothing.

The test drives it through the C ABI with caller-owned device buffers, twice in a row with
different problem sizes (the second call finds the first call's leftovers in the shared
scratch) and a third time on a stream of the caller's own, and compares with the regular entry
point, which allocates its scratch through the handle, and with the CPU reference's gradient.  From
the second call on it also checks that the free device memory did not shrink across the call.  It also checks that a problem too large for the fixed array is refused, not run.
This file sorts last on purpose: if it fails under `-x`, every other test has already run.
"""
import ctypes as C

import numpy as np
import pytest

from tests.util import RTOL_GRAD, RTOL_SCALAR, rel_grad, rel_sum

pytestmark = pytest.mark.gpu


def _static_call(lib, torch, dev, coeffs, t, diag, y, Jr, Jc, stream=None):
    P = 1 + 2 * Jr + 4 * Jc
    N = t.shape[0]
    to = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    dc, dt, dd, dy = to(coeffs), to(t), to(diag), to(y)
    val = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
    grad = torch.full((P,), float("nan"), dtype=torch.float64, device=dev)
    status = torch.full((1,), -1, dtype=torch.int32, device=dev)
    p = lambda a: C.c_void_p(a.data_ptr())  # noqa: E731
    torch.cuda.synchronize(dev)          # the inputs were written on torch's stream
    handle = torch.cuda.current_stream(dev).cuda_stream if stream is None else stream.cuda_stream  # 0 = default stream
    free_before = torch.cuda.mem_get_info(dev)[0]
    rc = lib.cel_grad_log_likelihood_static_scratch(N, Jr, Jc, p(dc), p(dt), p(dd), p(dy),
                                                    p(val), p(grad), p(status), C.c_void_p(handle))
    torch.cuda.synchronize(dev)
    free_after = torch.cuda.mem_get_info(dev)[0]
    return rc, float(val.item()), grad.cpu().numpy(), int(status.item()), free_before - free_after


def test_static_scratch_gradient_matches_regular_entry_point():
    import torch
    import oracle
    from celerite_b200 import _capi, synth
    from celerite_b200.batched import BatchedSolver
    lib = _capi.lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    regular = BatchedSolver(0)
    chk = oracle.best()
    side = torch.cuda.Stream(dev)        # the third call runs on a stream of the caller's own
    # first a long, wide problem, then a short, narrow one that reuses the same scratch words, then the
    # first again on a non-default stream
    for k, (N, Jr, Jc, seed) in enumerate([(6000, 1, 3, 11), (257, 0, 1, 12), (6000, 1, 3, 11)]):
        b = synth.make_batch(1, N, Jr, Jc, seed=seed)
        rc, val, grad, status, grown = _static_call(lib, torch, dev, b["coeffs"][0], b["t"][0], b["diag"][0],
                                                    b["y"][0], Jr, Jc, stream=side if k == 2 else None)
        assert rc == _capi.OK and status == 0
        # the entry point that allocates its scratch through a handle: same kernel, other memory
        ref_val, ref_grad, ref_status = regular.grad_log_likelihood(b["coeffs"], b["t"], b["diag"], b["y"], Jr, Jc)
        assert ref_status[0] == 0
        assert rel_sum(val, ref_val[0], N) < RTOL_SCALAR
        assert rel_grad(grad, ref_grad[0]) < RTOL_GRAD
        # and the CPU reference's forward-mode gradient (solver.cpp:347-463; its value carries -0.5*pi*log N)
        j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], Jr, Jc)
        st, cpu_val, cpu_grad = chk.grad(oracle.Problem(j, ar, cr, ac, bc, cc, dc, b["t"][0], b["diag"][0]), b["y"][0])
        assert st == 0
        assert rel_sum(val - 0.5 * np.pi * np.log(N), cpu_val, N) < RTOL_SCALAR
        assert rel_grad(grad, cpu_grad) < RTOL_GRAD
        # after the first call on the device (which may load the library's device image), a call must
        # leave the amount of free device memory where it was
        if k > 0:
            assert grown <= 0, "device memory shrank by %d bytes across an allocation-free call" % grown


def test_static_scratch_refuses_what_does_not_fit():
    import torch
    from celerite_b200 import _capi
    lib = _capi.lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    # J = 32 at N = 3e7 needs ~52 MiB of checkpoints + tape: more than the 32 MiB array.  The call
    # must say so before touching any of its arguments, so small dummy buffers are enough.
    dummy = torch.zeros(128, dtype=torch.float64, device=dev)
    st = torch.zeros(1, dtype=torch.int32, device=dev)
    p = lambda a: C.c_void_p(a.data_ptr())  # noqa: E731
    rc = lib.cel_grad_log_likelihood_static_scratch(30_000_000, 0, 16, p(dummy), p(dummy), p(dummy), p(dummy),
                                                    p(dummy), p(dummy), p(st), None)
    assert rc == _capi.EXCEEDS_STATIC_SCRATCH
    rc = lib.cel_grad_log_likelihood_static_scratch(100, 0, 17, p(dummy), p(dummy), p(dummy), p(dummy),
                                                    p(dummy), p(dummy), p(st), None)
    assert rc == _capi.UNSUPPORTED      # J = 34 > 32
