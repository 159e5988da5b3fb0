"""Data-parallel training / sharded retrieval on real GPUs (-m gpu): tests/dp_check.py under torchrun.
With >= 2 visible GPUs: one rank per GPU over NCCL.  On ANY box, a single-GPU one included: two ranks as two processes
on cuda:0 -- the peer-memory path (clsm_dp_push_rows / clsm_adam_table_dp over CUDA IPC mappings, share_across_ranks,
the ordering collectives) runs exactly as on two GPUs, only the small collectives go over gloo."""
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


def _run(env_extra, nproc):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ)
    env.update(env_extra)
    return subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
                           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "dp_check.py")],
                          capture_output=True, text=True, timeout=900, env=env)


def test_two_ranks_on_one_gpu_match_single_process():
    """N ranks on 1/N of a batch == one process on the whole batch (first-step gradients, loss trajectory, weights,
    identical replicas, sharded retrieval), for fp32 (NCCL-style exchange path) and bf16 (peer-memory kernels)."""
    out = _run({"DP_CHECK_ONE_GPU": "1"}, 2)
    assert "DP_CHECK_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
