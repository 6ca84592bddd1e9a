"""CPU only: run the literal oracle over the first P pods of a config at its full node count and store the compact
per-pod outcome (status, host, filtered count, score, FitError text hash) for a later comparison with the GPU engine
(scripts/compare_prefix.py). Needs no GPU, so it can run wherever there are spare cores."""
import hashlib
import json
import sys
import time

sys.path.insert(0, ".")
from kubernetes_b200 import workloads
from tests import oracle_lib

cfg, pods_prefix, workers, out_path = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
nodes, pods = workloads.CONFIG_SIZES[cfg]
cluster = workloads.generate(cfg, nodes, pods)
cluster["pods"] = cluster["pods"][:pods_prefix]
t0 = time.time()
want = oracle_lib.call({"op": "stream", "cluster": cluster, "workers": workers})
assert not want.get("error"), want.get("error")
rows = [[r["status"], r["host"], r["filtered"], r["score"], hashlib.sha1(r["message"].encode()).hexdigest()[:12] if r["status"] == 2 else ""]
        for r in want["results"]]
json.dump({"config": cfg, "nodes": nodes, "stream_pods": pods, "prefix": pods_prefix, "oracle_seconds": round(time.time() - t0, 1),
           "workers": workers, "lastNodeIndex": want["lastNodeIndex"], "results": rows}, open(out_path, "w"), separators=(",", ":"))
print("done", time.time() - t0)
