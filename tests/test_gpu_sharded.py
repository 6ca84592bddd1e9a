"""Sharded node table on two GPUs: placements must equal the unsharded run bit for bit (scripts/shard_check.py
compares placement hashes and the round-robin counter on every rank). Skipped when fewer than two GPUs."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("cfg,nodes,pods", [(3, 1500, 1500), (5, 1200, 1200), (2, 600, 1500)])
def test_two_shards_match_one(cfg, nodes, pods):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29641", os.path.join("scripts", "shard_check.py"), str(cfg), str(nodes), str(pods)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert '"match": true' in r.stdout
