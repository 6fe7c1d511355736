"""GPU, 2+ devices: the library's NCCL path (mfb_comm_*, distributed mfb_pcg_dev / mfb_newton_step) under torchrun,
checked against the single-GPU result of the same global problem.  Skipped on a 1-GPU box; the host-side partition
logic is covered on CPU with gloo in tests/test_partition_gloo.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count()


def _torchrun(script, nproc, port, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "dist", script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    assert r.returncode == 0, (r.stdout[-3000:], r.stderr[-3000:])
    assert r.stdout.count("DIST_OK") == nproc, r.stdout[-3000:]
    return r.stdout


def test_distributed_pcg_matches_single_gpu(built):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    _torchrun("run_dist_pcg.py", 2, 29531)


def test_distributed_newton_step_matches_single_gpu(built):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    _torchrun("run_dist_newton.py", 2, 29533)


def test_distributed_update_with_overlapped_exchange_matches_single_gpu(built):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    _torchrun("run_dist_update.py", 2, 29535)
