"""Multi-GPU parity of the end-of-denoise exchange: needs >= 2 GPUs on the box, skipped otherwise.
(The N > 1 host logic is covered on CPU by test_gloo_world2.py.)"""
import os
import subprocess
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.gpu
def test_peer_memory_blend_equals_nccl_and_single():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29531", os.path.join(HERE, "_peer_blend_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "PEER_BLEND_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.gpu
def test_halo_exchange_two_gpus_equals_single_process():
    """per-step halo exchange over NVLink peer memory == the same plan in one process, bit for bit."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(HERE, "_halo_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "HALO_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
