"""Sweep the speculative group size on one synthetic config; prints kernel time, groups, restarts."""
import json
import sys

sys.path.insert(0, ".")
from kubernetes_b200 import lib, workloads

cfg, nodes, pods = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
sizes = [int(x) for x in sys.argv[4].split(",")] if len(sys.argv) > 4 else [1, 2, 4, 8, 16, 32]
pipeline = int(sys.argv[5]) if len(sys.argv) > 5 else 0
cluster = workloads.generate(cfg, nodes, pods)
sess = lib.Session(cluster=cluster)
sess.run(results=False)
for g in sizes:
    out = sess.run(results=False, group_size=g, replay_iters=2, pipeline=pipeline)
    print(json.dumps({"cfg": cfg, "nodes": nodes, "pods": pods, "G": g, "pipeline": pipeline, "kernel_ms": round(out["kernel_ms"], 2),
                      "replay_ms": round(out["replay_ms"], 2), "us_per_pod": round(1000 * out["replay_ms"] / pods, 2),
                      "groups": out["groups"], "restarts": out["group_restarts"], "encode_ms": round(out["encode_ms"], 1),
                      "e2e_ms": round(out["e2e_ms"], 1), "decode_ms": round(out["decode_ms"], 1), "post_ms": round(out["post_ms"], 1), "sync_ms": round(out["sync_ms"], 1), "hash": out["placement_hash"][-6:],
                      "raw_us_per_pod": [round(c / 1965.0 / pods, 3) for c in out["stage_cycles"][:12]],
                      "stage_us_per_pod": [round(c / 1965.0 / pods, 3) for c in out["stage_cycles"][:7]],
                      "p3": {"rounds": out["stage_cycles"][7], "window_pods": out["stage_cycles"][8], "one_pod_path": out["stage_cycles"][9],
                             "setup_us_per_pod": round(out["stage_cycles"][10] / 1965.0 / pods, 3),
                             "decide_us_per_pod": round(out["stage_cycles"][11] / 1965.0 / pods, 3)}}))
sess.close()
