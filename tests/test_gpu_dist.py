"""Row-partitioned mode on real GPUs (needs >= 2): the partitioned solve must reproduce the
single-GPU solve of the same QP (status, objective, solution) -- tools/dist_check.py."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_partitioned_matches_single_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29533",
           os.path.join(ROOT, "tools", "dist_check.py"), "3000", "1e-4"]
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=550)
    assert r.returncode == 0, r.stdout[-3000:]
    assert r.stdout.count("ok=True") == 2, r.stdout[-3000:]
