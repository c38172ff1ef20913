"""GPU, 2 ranks over NCCL: the sharded paths on real devices (self-skips below 2 GPUs).  Launches
tools/sharded_check.py under torch.distributed.run: candidate-sharded plans (fp32 / f16+refine / f16x3; one-shot and
PlanRing) must equal the single-GPU plan over the same K_total bit for bit, and the row-sharded ndiv loss/gradient
(ShardedNdiv: two all-gathers on libndivplan's own communicator) must equal the single-GPU call."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sharded_paths_two_ranks_nccl(cuda):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tools", "sharded_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-4000:]
    assert "sharded_check ok" in r.stdout
