"""Data-parallel training step == single-process step (VERDICT r1 missing #3): BatchNorm (sum x, sum x^2, n) /
(sum dy, sum dy xhat) all-reduce through the library's hook (cm_set_bn_sync / training.BatchNormSync) + gradient SUM
all-reduce.  Two ranks; on a one-GPU box both share cuda:0 over gloo, with two GPUs they use NCCL."""
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


def _run(backend):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "dist_train_check.py"), "--backend", backend]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "DIST_TRAIN_OK" in r.stdout, (r.stdout[-2000:], r.stderr[-3000:])


def test_two_ranks_equal_single_process_gloo_shared_gpu():
    _run("gloo")


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_ranks_equal_single_process_nccl():
    _run("nccl")
