"""Multi-rank parity of the PRODUCT code (row-sharded dense solves, source-sharded empirical solve, problem-sharded
batched solve, NVLink peer exchange): spawns tests/mgpu_worker.py under torchrun with 2 ranks.

  test_two_gpu_parity        : one rank per GPU over NCCL (needs 2 GPUs; skipped on a single-GPU box);
  test_two_ranks_one_gpu     : both ranks on cuda:0 (runs on ANY GPU box): gloo for the host plumbing, the peer
                               exchange buffers CUDA-IPC-mapped between the two processes on the same device - the
                               same pot_b200.sharded / comm kernels as the multi-GPU run, time-sliced on one GPU."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpu_parity():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "mgpu_worker.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    sys.stdout.write(r.stdout[-4000:])
    sys.stderr.write(r.stderr[-4000:])
    assert r.returncode == 0 and "MGPU_OK" in r.stdout


def test_two_ranks_one_gpu():
    env = dict(os.environ, MGPU_ONE_DEVICE="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "mgpu_worker.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900, env=env)
    sys.stdout.write(r.stdout[-4000:])
    sys.stderr.write(r.stderr[-4000:])
    assert r.returncode == 0 and "MGPU_OK" in r.stdout
