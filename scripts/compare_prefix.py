"""GPU side of scripts/oracle_prefix.py: schedule the same prefix of the stream on the engine and compare every pod
with the stored oracle outcome. Prints one JSON line."""
import hashlib
import json
import sys

sys.path.insert(0, ".")
from kubernetes_b200 import lib, workloads

ref = json.load(open(sys.argv[1]))
cfg, prefix = ref["config"], ref["prefix"]
cluster = workloads.generate(cfg, ref["nodes"], ref["stream_pods"])
cluster["pods"] = cluster["pods"][:prefix]
sess = lib.Session(cluster=cluster)
got = sess.run()
sess.close()
bad = []
for i, (g, w) in enumerate(zip(got["results"], ref["results"])):
    msg = hashlib.sha1(g["message"].encode()).hexdigest()[:12] if g["status"] == 2 else ""
    if [g["status"], g["host"], g["filtered"], g["score"] if g["status"] == 0 else w[3], msg] != w:
        bad.append(i)
print(json.dumps({"config": cfg, "oracle_mode": "literal", "nodes": ref["nodes"], "pods": prefix,
                  "note": "first %d pods of the %d-pod stream at the full node count" % (prefix, ref["stream_pods"]),
                  "compared": len(ref["results"]), "mismatches": len(bad), "first_mismatches": bad[:5],
                  "last_node_index_equal": got["lastNodeIndex"] == ref["lastNodeIndex"], "scheduled": got["scheduled"],
                  "fit_errors": got["fit_errors"], "oracle_seconds": ref["oracle_seconds"], "gpu_e2e_ms": round(got["e2e_ms"], 1),
                  "placement_hash": got["placement_hash"]}))
