"""Row-sharded recall over NCCL on real GPUs (needs >= 2 devices; skipped on a 1-GPU box, where
tests/test_gpu_merge.py covers the same merge with G stores on one device and tests/test_dist_gloo.py
covers the plumbing)."""
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.parametrize("exchange", ["nccl", "p2p"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_recall_matches_oracle(world, exchange):
    """exchange = nccl: host-issued allgather + merge; p2p: the library's fused exchange over peer memory."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import os
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + world + (10 if exchange == "p2p" else 0)),
           str(ROOT / "tests" / "dist_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, EXCHANGE=exchange))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count(": ok") == 3
