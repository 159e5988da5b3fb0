"""Data-parallel training / sharded retrieval on real GPUs (-m gpu).  Needs >= 2 visible GPUs: launches
tests/dp_check.py under torchrun (NCCL); skipped on a single-GPU box, where tests/test_parallel_gloo.py
covers the host plumbing."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_rank_training_matches_single_process():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "dp_check.py")],
                         capture_output=True, text=True, timeout=900)
    assert "DP_CHECK_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
