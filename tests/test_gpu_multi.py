"""Two GPUs: one long series split by time, span carries exchanged by NCCL inside the C ABI
(cel_comm_init, cel_compute_mgpu, cel_solve_mgpu, cel_dot_L_mgpu, cel_log_likelihood_mgpu) and, for
comparison, through torch.distributed; both must reproduce one GPU holding the whole series to 1e-10
(tools/dist_check.py does the work, one process per GPU under torchrun).  Skipped on a one-GPU box."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("N", [400_000, 1_000_003])
def test_long_series_over_two_gpus(N):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tools", "dist_check.py"), str(N)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "OK" in r.stdout, (r.stdout + r.stderr)[-3000:]
    assert "inside the C ABI" in r.stdout


def test_cpp_program_over_the_c_abi_alone():
    """examples/mgpu_long_series.cu: one host thread per visible GPU, no Python and no PyTorch in the
    process -- cel_comm_unique_id / cel_comm_init / cel_compute_mgpu / cel_solve_mgpu / cel_dot_L_mgpu /
    cel_log_likelihood_mgpu against one GPU holding the whole series (with one GPU the world is 1 and
    the collectives degenerate; with more, NCCL carries the span summaries)."""
    exe = os.path.join(ROOT, "build", "bin", "mgpu_long_series")
    if not os.path.exists(exe):
        pytest.skip("build/bin/mgpu_long_series is not built (python -m celerite_b200.build)")
    r = subprocess.run([exe, "600000"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "OK" in r.stdout, (r.stdout + r.stderr)[-2000:]
