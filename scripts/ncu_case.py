"""One pass of a synthetic config through ksb_schedule_batch, sized to stay a single kernel launch (< 8,192 pods):
the command profiled by ncu (profiles/README.md). usage: ncu_case.py CFG NODES PODS [pipeline]"""
import json
import sys

sys.path.insert(0, ".")
from kubernetes_b200 import lib, workloads

cfg, nodes, pods = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
opts = {"results": False}
if len(sys.argv) > 4:
    opts["pipeline"] = int(sys.argv[4])
cluster = workloads.generate(cfg, nodes, pods)
sess = lib.Session(cluster=cluster)
out = sess.run(**opts)
sess.close()
print(json.dumps({"cfg": cfg, "nodes": nodes, "pods": pods, "kernel_ms": out["kernel_ms"], "launches": out["kernel_launches"],
                  "groups": out["groups"], "hash": out["placement_hash"]}))
