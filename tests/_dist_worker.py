"""Worker for tests/test_multi_rank_cpu.py: world_size-2 gloo process on CPU. Checks the rank plumbing that
bench.py and the sharded engine rely on (row plan, handle exchange by all_gather_object, max-over-ranks
timing reduction) without touching a GPU."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kubernetes_b200 import sharding, workloads

rank = int(os.environ["RANK"])
world = int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
n_nodes = int(sys.argv[1])
cluster = workloads.generate(3, n_nodes, 50)
# every rank builds the same stream (replicated host state)
names = [n["name"] for n in cluster["nodes"]]
digest = [None] * world
dist.all_gather_object(digest, (len(names), names[0], names[-1], cluster["pods"][7]["name"]))
assert all(d == digest[0] for d in digest), digest
# the row plan partitions [0, n) in order
plan = sharding.plan(n_nodes, world)
mine = plan[rank]
allp = [None] * world
dist.all_gather_object(allp, mine)
pos = 0
for base, cnt in allp:
    assert base == pos or cnt == 0, allp
    pos = base + cnt
assert pos == n_nodes, allp
assert all(b % 256 == 0 for b, c in allp if c), allp


class FakeEngine:  # stands in for lib.OpenEngine: the exchange logic is what is under test
    def __init__(self):
        self.connected = None

    def shard_export(self, max_batch_pods):
        return b"blob-of-rank-%d-%d" % (rank, max_batch_pods)

    def shard_connect(self, blobs):
        self.connected = blobs


e = FakeEngine()
sharding.connect(e, rank, world, dist.all_gather_object, 123)
assert e.connected == [b"blob-of-rank-%d-123" % r for r in range(world)], e.connected
# bench.py's reduction: device time is the max over ranks
t = torch.tensor([10.0 + rank, 5.0 - rank], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
assert t.tolist() == [10.0 + world - 1, 5.0], t
dist.barrier()
dist.destroy_process_group()
print("rank %d ok" % rank)
