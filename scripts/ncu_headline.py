"""The headline workload for `ncu`: cfg4 (50,000 nodes x 100,000 pods) once end to end (26 chunked launches of the
batch kernel) and then one resident replay of the whole batch in a single launch, which is the launch bench.py
times as `value`. Capture that last launch with `-k regex:ksb_schedule_batch -s 25 -c 1` (profiles/README.md)."""
import json
import sys

sys.path.insert(0, ".")
from kubernetes_b200 import lib, workloads

cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 4
n_nodes, n_pods = workloads.CONFIG_SIZES[cfg]
cluster = workloads.generate(cfg, n_nodes, n_pods)
sess = lib.Session(cluster=cluster)
out = sess.run(results=False, replay_iters=1)
sess.close()
print(json.dumps({"cfg": cfg, "nodes": n_nodes, "pods": n_pods, "launches": out["kernel_launches"], "replay_ms": out["replay_ms"],
                  "hash": out["placement_hash"]}))
