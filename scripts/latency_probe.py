"""One-pod-per-call latency (ksb_schedule + AssumePod): percentiles and the slowest calls of a stream prefix."""
import json
import sys

sys.path.insert(0, ".")
from kubernetes_b200 import lib, workloads

cfg, nodes, pods, limit = (int(x) for x in sys.argv[1:5])
sess = lib.Session(cluster=workloads.generate(cfg, nodes, pods))
for _ in range(2):
    out = sess.run(results=False, batch=0, limit=limit)
    print(json.dumps({k: out[k] for k in ("p50_us", "p99_us", "max_us", "slowest")}))
sess.close()
