"""Sharded objectives, psislw and divergence bound over NCCL against the unsharded oracle: one process per GPU
(scripts/check_multi_gpu.py under torchrun).  Needs at least two visible GPUs; skipped on a single-GPU box."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
def test_two_rank_nccl_parity():
    world = 2
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(world),
           '--master-addr', '127.0.0.1', '--master-port', '29533', os.path.join(ROOT, 'scripts', 'check_multi_gpu.py')]
    res = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert 'MULTI_GPU_OK world=%d' % world in res.stdout
