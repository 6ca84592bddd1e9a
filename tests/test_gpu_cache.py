"""The scheduler cache's pod state machine (schedulercache/cache.go:118-312, expiry :426-462) over the GPU engine:
the host-side CacheDecorator keeps the states, the engine the node aggregates. Checked against the vectors
transcribed from cache_test.go and, on random operation sequences, against the oracle's restatement. ksb_node_info
also compares the device columns of each reported row with the host records."""
import json
import os

import pytest

from kubernetes_b200 import lib, workloads
from tests import oracle_lib

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_golden_cache_vectors_on_gpu():
    with open(os.path.join(HERE, "golden", "reference_vectors.json")) as f:
        cases = [c for c in json.load(f)["cases"] if c["op"] == "cache_ops"]
    assert len(cases) >= 12
    for c in cases:
        req = {k: v for k, v in c.items() if k not in ("expect", "source", "test", "op")}
        out = lib.cache_ops(req)
        assert out["errors"] == c["expect"]["errors"], (c["test"], out["errors"])
        assert out["nodes"] == c["expect"]["nodes"], (c["test"], out["nodes"])


def _lcg(seed):
    state = seed
    while True:
        state = (state * 6364136223846793005 + 1442695040888963407) % (1 << 64)
        yield state >> 33


@pytest.mark.parametrize("cfg,seed", [(3, 1), (3, 2), (2, 3), (5, 4)])
def test_random_cache_sequences_match_oracle(cfg, seed):
    """Random walks over the state machine, valid and invalid transitions alike, on nodes with and without a Node
    object; errors and the aggregates of every node must equal the oracle's."""
    cluster = workloads.generate(cfg, 12, 60)
    nodes = cluster["nodes"]
    names = [n["name"] for n in nodes] + ["ghost-node"]  # pods may sit on a node the cache has no Node object for
    rnd = _lcg(seed)
    ops = [{"do": "add_node", "node": n} for n in nodes[:10]]
    where = {}
    now = 0.0
    for _ in range(400):
        r = next(rnd) % 100
        k = next(rnd) % len(cluster["pods"])
        pod = dict(cluster["pods"][k])
        pod.pop("extendedResources", None)  # device assignment needs the bind echo; aggregates are what is checked here
        pod["nodeName"] = where.get(k) or names[next(rnd) % len(names)]
        now += (next(rnd) % 40) / 10.0
        if r < 30:
            ops += [{"do": "assume", "pod": pod}, {"do": "finish", "pod": pod, "now": now}]
            where.setdefault(k, pod["nodeName"])
        elif r < 50:
            ops.append({"do": "add", "pod": pod})
            where.setdefault(k, pod["nodeName"])
        elif r < 60:
            ops.append({"do": "forget", "pod": pod})
        elif r < 72:
            ops.append({"do": "remove", "pod": pod})
        elif r < 82:
            # An update changes labels only. Container resources and ports of a pod are immutable in the API, and the
            # reference relies on it: UpdatePod leaves the cached object of podStates untouched and RemovePod later
            # subtracts that stale object's resources (cache.go:259-312), so an update that changed them would leave
            # a residue in the reference that the engine, which subtracts what it stored, does not reproduce.
            new = dict(pod)
            new["labels"] = dict(pod.get("labels", {}), rev="r%d" % (next(rnd) % 5))
            ops.append({"do": "update", "old": pod, "pod": new})
        elif r < 92:
            ops.append({"do": "cleanup", "now": now})
        elif r < 96:
            ops.append({"do": "is_assumed", "pod": pod})
        elif r < 98:
            ops.append({"do": "add_node", "node": nodes[next(rnd) % len(nodes)]})
        else:
            ops.append({"do": "remove_node", "name": names[next(rnd) % len(names)]})
    req = {"ttl": 5, "ops": ops, "report": names}
    want = oracle_lib.call(dict(req, op="cache_ops"))
    assert not want.get("error"), want
    got = lib.cache_ops(req)
    assert got["errors"] == want["errors"], [(i, ops[i]["do"], g, w) for i, (g, w) in enumerate(zip(got["errors"], want["errors"])) if g != w][:5]
    assert got["nodes"] == want["nodes"], {n: (got["nodes"][n], want["nodes"][n]) for n in names if got["nodes"][n] != want["nodes"][n]}
