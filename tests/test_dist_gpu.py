"""Two-rank parity of the fused training loop's in-kernel all-reduce (NVLink peer memory) against
the step-wise NCCL path; needs two GPUs on the box, skipped otherwise."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
def test_fused_loop_two_ranks_match_stepwise_nccl():
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2',
           '--master-addr', '127.0.0.1', '--master-port', '29531', os.path.join(ROOT, 'tools', 'dist_check.py')]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert 'DIST_CHECK OK' in out.stdout, out.stdout[-2000:]
