"""Full-size parity run (not part of the test suite: the literal oracle needs minutes at these sizes).
For each config given, the whole stream goes through the GPU engine and through the CPU oracle, and every pod's
status, host, filtered count, score, device ids and FitError text are compared. Prints one JSON line per config."""
import json
import sys
import time

sys.path.insert(0, ".")
from kubernetes_b200 import lib, workloads
from tests import oracle_lib

for spec in sys.argv[1:] or ["1", "2", "3"]:  # "cfg" (BASELINE.json size) or "cfg:nodes:pods"
    indexed = spec.endswith("i")  # "5i": the oracle's indexed mode (inter-pod affinity without the O(nodes x pods) rescans)
    parts = [int(x) for x in spec.rstrip("i").split(":")]
    cfg = parts[0]
    nodes, pods = parts[1:3] if len(parts) == 3 else workloads.CONFIG_SIZES[cfg]
    cluster = workloads.generate(cfg, nodes, pods)
    t0 = time.time()
    want = oracle_lib.call({"op": "stream", "cluster": cluster, "workers": 16, "indexed": indexed})
    t_oracle = time.time() - t0
    assert not want.get("error"), want.get("error")
    sess = lib.Session(cluster=cluster)
    got = sess.run()
    sess.close()
    bad = []
    for i, (g, w) in enumerate(zip(got["results"], want["results"])):
        same = (g["status"], g["host"], g["filtered"]) == (w["status"], w["host"], w["filtered"])
        if same and w["status"] == 0:
            same = g["score"] == w["score"] and g["devices"] == w["devices"]
        if same and w["status"] == 2:
            same = g["message"] == w["message"]
        if not same:
            bad.append(i)
    print(json.dumps({"config": cfg, "oracle_mode": "indexed" if indexed else "literal", "nodes": nodes, "pods": pods, "compared": len(want["results"]),
                      "mismatches": len(bad), "first_mismatches": bad[:5],
                      "last_node_index_equal": got["lastNodeIndex"] == want["lastNodeIndex"],
                      "scheduled": got["scheduled"], "fit_errors": got["fit_errors"],
                      "oracle_seconds": round(t_oracle, 1), "gpu_e2e_ms": round(got["e2e_ms"], 1),
                      "placement_hash": got["placement_hash"]}), flush=True)
