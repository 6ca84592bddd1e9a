"""Race hunt: the same stream many times through the pipelined schedule; every run must give the plain schedule's
placements. usage: repeat_check.py CFG NODES PODS RUNS [group sizes, comma separated]"""
import json
import sys

sys.path.insert(0, ".")
from kubernetes_b200 import lib, workloads

cfg, nodes, pods, runs = (int(x) for x in sys.argv[1:5])
sizes = [int(x) for x in sys.argv[5].split(",")] if len(sys.argv) > 5 else [4, 16]
pipeline = int(sys.argv[6]) if len(sys.argv) > 6 else 1
cluster = workloads.generate(cfg, nodes, pods)
sess = lib.Session(cluster=cluster)
base = sess.run(results=True, pipeline=-1, group_size=32)
bad = []
for i in range(runs):
    for g in sizes:
        out = sess.run(results=True, pipeline=pipeline, group_size=g)
        if out["placement_hash"] != base["placement_hash"]:
            first = next(j for j, (a, b) in enumerate(zip(out["results"], base["results"])) if (a["status"], a["host"]) != (b["status"], b["host"]))
            bad.append({"run": i, "G": g, "first_pod": first, "got": out["results"][first], "want": base["results"][first], "groups": out["groups"]})
sess.close()
print(json.dumps({"cfg": cfg, "nodes": nodes, "pods": pods, "runs": runs * len(sizes), "mismatching_runs": len(bad), "first": bad[:3]}))
