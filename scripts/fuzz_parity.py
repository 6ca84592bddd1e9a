"""Seeded sweep: GPU engine vs literal oracle on many small random clusters, alternating the plain and the pipelined
schedule and several group sizes. Prints one line per case and a summary; exits 1 on the first mismatch."""
import json
import sys

sys.path.insert(0, ".")
from kubernetes_b200 import lib, workloads
from tests import oracle_lib

n_seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 8
cases = 0
for seed in range(100, 100 + n_seeds):
    for cfg, nodes, pods in ((3, 150 + 37 * (seed % 5), 900), (5, 60 + 11 * (seed % 7), 500), (2, 40 + seed % 30, 500), (1, 20 + seed % 40, 600)):
        cluster = workloads.generate(cfg, nodes, pods, seed=seed * 7919 + cfg)
        want = oracle_lib.call({"op": "stream", "cluster": cluster, "workers": 8})
        assert not want.get("error"), want.get("error")
        sess = lib.Session(cluster=cluster)
        for opts in ({"pipeline": -1, "group_size": 32}, {"pipeline": 1, "group_size": 16}, {"pipeline": 1, "group_size": 2 + seed % 9},
                     {"pipeline": -1, "group_size": 1 + seed % 7}, {"batch": 0}):
            got = sess.run(**opts)
            for i, (g, w) in enumerate(zip(got["results"], want["results"])):
                same = (g["status"], g["host"], g["filtered"]) == (w["status"], w["host"], w["filtered"])
                if same and w["status"] == 0:
                    same = g["score"] == w["score"] and g["devices"] == w["devices"]
                if same and w["status"] == 2:
                    same = g["message"] == w["message"]
                if not same:
                    print(json.dumps({"mismatch": True, "seed": seed, "cfg": cfg, "nodes": nodes, "opts": opts, "pod": i, "gpu": g, "oracle": w}))
                    sys.exit(1)
            assert got["lastNodeIndex"] == want["lastNodeIndex"], (seed, cfg, opts)
            cases += 1
        sess.close()
    print(json.dumps({"seed": seed, "cases_so_far": cases}), flush=True)
print(json.dumps({"cases": cases, "mismatches": 0}))
