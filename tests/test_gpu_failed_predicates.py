"""ksb_failed_predicates: the per-node failure reasons of one Schedule call (the reference's FailedPredicateMap,
core/generic_scheduler.go:283-362) must equal, node by node, what the oracle's podFitsOnNode reports."""
import pytest

from kubernetes_b200 import lib, workloads
from tests import oracle_lib

pytestmark = pytest.mark.gpu


def _oracle_failures(cluster, pod):
    out = {}
    for node in cluster["nodes"]:
        r = oracle_lib.call({"op": "predicate", "name": "all", "cluster": cluster, "pod": pod, "node": node["name"]})
        assert not r.get("error"), r
        if r["reasons"]:
            out[node["name"]] = sorted(set(r["reasons"]))
        elif pod.get("extendedResources"):
            # the fork's device check runs on the nodes that passed every predicate (generic_scheduler.go:357)
            r = oracle_lib.call({"op": "predicate", "name": "HasExtendedResources", "cluster": cluster, "pod": pod,
                                 "node": node["name"]})
            assert not r.get("error"), r
            if r["reasons"]:
                out[node["name"]] = sorted(set(r["reasons"]))
    return out


@pytest.mark.parametrize("cfg,nodes,pods,picks", [(3, 120, 400, (0, 3, 17, 58, 131, 399)), (2, 60, 300, (0, 5, 250)),
                                                  (5, 90, 300, (0, 1, 40))])
def test_failed_predicates_match_oracle(cfg, nodes, pods, picks):
    base = workloads.generate(cfg, nodes, pods)
    # a partly filled cluster: the first half of the stream is placed by the engine and becomes existing pods
    sess = lib.Session(cluster=base)
    placed = sess.run(limit=pods // 2)
    sess.close()
    existing = []
    for p, r in zip(base["pods"], placed["results"]):
        if r["status"] == 0:
            q = dict(p)
            q["nodeName"] = r["host"]
            if r["devices"]:
                q["extendedResources"] = [dict(e, assigned=r["devices"].get(e["name"], [])) for e in p.get("extendedResources", [])]
            existing.append(q)
    checked = 0
    for k in picks:
        pod = base["pods"][k]
        cluster = dict(base, existingPods=existing, pods=[pod])
        s2 = lib.Session(cluster=cluster)
        got = s2.run(batch=0, failed_predicates=True)
        s2.close()
        res = got["results"][0]
        want = _oracle_failures(cluster, pod)
        have = {n: sorted(set(v)) for n, v in res.get("failed", {}).items()}
        assert have == want, (cfg, k, {n: (have.get(n), want.get(n)) for n in set(have) | set(want) if have.get(n) != want.get(n)})
        assert res["filtered"] == nodes - len(want)
        checked += len(want)
    assert checked > 0
