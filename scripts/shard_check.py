"""torchrun -N ranks: schedule one synthetic stream with the node table sharded over the ranks' GPUs and
check the placements against an unsharded run on rank 0's GPU. Prints timing per rank."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, ".")
from kubernetes_b200 import lib, workloads

cfg, nodes, pods = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
gsize = int(sys.argv[4]) if len(sys.argv) > 4 else 0
pipeline = int(sys.argv[5]) if len(sys.argv) > 5 else 0
rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("gloo")
cluster = workloads.generate(cfg, nodes, pods)
sess = lib.Session(cluster=cluster)
ref = sess.run(device=local, results=False) if rank == 0 else None
eng = lib.OpenEngine(sess, device=local, shard_rank=rank, shard_count=world, group_size=gsize, pipeline=pipeline)
blob = eng.shard_export(pods)
blobs = [None] * world
dist.all_gather_object(blobs, blob)
eng.shard_connect(blobs)
dist.barrier()
out = eng.run(results=False, replay_iters=2)
res = {"rank": rank, "hash": out["placement_hash"], "kernel_ms": round(out["kernel_ms"], 2),
       "replay_ms": round(out["replay_ms"], 2), "groups": out["groups"], "restarts": out["group_restarts"],
       "scheduled": out["scheduled"], "lastNodeIndex": out["lastNodeIndex"],
       "stage_us_per_pod": [round(c / 1965.0 / pods, 2) for c in out["stage_cycles"]]}
all_res = [None] * world
dist.all_gather_object(all_res, res)
if rank == 0:
    for r in all_res:
        print(json.dumps(r))
    ok = all(r["hash"] == ref["placement_hash"] for r in all_res) and all(
        r["lastNodeIndex"] == ref["lastNodeIndex"] for r in all_res)
    print(json.dumps({"unsharded_hash": ref["placement_hash"], "unsharded_kernel_ms": round(ref["kernel_ms"], 2),
                      "match": ok}))
    if not ok:
        sys.exit(1)
eng.close()
dist.destroy_process_group()
