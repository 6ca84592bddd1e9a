"""What the Policy instantiation of the kernel costs next to the DefaultProvider one: resident replay time of the same
stream with and without a Policy-only priority enabled. python scripts/policy_cost.py [CFG]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kubernetes_b200 import lib, workloads  # noqa: E402

cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 3
nodes, pods = {3: (5000, 30000), 4: (50000, 100000), 5: (5000, 20000)}[cfg]
sess = lib.Session(cluster=workloads.generate(cfg, nodes, pods))
base = {"LeastRequestedPriority": 1, "BalancedResourceAllocation": 1, "SelectorSpreadPriority": 1, "NodeAffinityPriority": 1,
        "TaintTolerationPriority": 1, "InterPodAffinityPriority": 1, "NodePreferAvoidPodsPriority": 10000}
out = {}
for name, extra in [("default", {}), ("equal", {"EqualPriority": 1}), ("most_requested", {"MostRequestedPriority": 1}),
                    ("default_again", {})]:
    r = sess.run(results=False, replay_iters=5, weights=dict(base, **extra))
    out[name] = {"replay_ms": r["replay_ms"], "pods_per_s": pods / (r["replay_ms"] / 1000.0)}
sess.close()
print(json.dumps({"cfg": cfg, "nodes": nodes, "pods": pods, "runs": out}))
