"""The oracle's indexed mode (one scan of the pods per Schedule instead of one per node) must give the literal mode's
results, field by field, before it is trusted at sizes the literal mode cannot reach."""
import json
import os

import pytest

from kubernetes_b200 import workloads
from tests import oracle_lib
from tests.test_oracle_golden import check_case

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("cfg,nodes,pods,seed", [(5, 60, 500, None), (5, 120, 700, 11), (5, 40, 600, 12), (3, 80, 400, None),
                                                 (2, 30, 200, None)])
def test_indexed_mode_equals_literal_mode(cfg, nodes, pods, seed):
    cluster = workloads.generate(cfg, nodes, pods, seed=seed)
    lit = oracle_lib.call({"op": "stream", "cluster": cluster, "workers": 4})
    idx = oracle_lib.call({"op": "stream", "cluster": cluster, "workers": 4, "indexed": True})
    assert not lit.get("error") and not idx.get("error")
    assert idx["lastNodeIndex"] == lit["lastNodeIndex"]
    assert len(idx["results"]) == len(lit["results"])
    for i, (a, b) in enumerate(zip(idx["results"], lit["results"])):
        assert a == b, (i, a, b)
    assert any(r["status"] == 0 for r in lit["results"])


def test_indexed_mode_on_the_golden_vectors():
    """Every transcribed stream / priority vector and the inter-pod affinity predicate vectors, in indexed mode."""
    with open(os.path.join(HERE, "golden", "reference_vectors.json")) as f:
        cases = [c for c in json.load(f)["cases"]
                 if c["op"] in ("stream", "priority") or c.get("name") in ("MatchInterPodAffinity", "all")]
    assert len(cases) > 60
    for c in cases:
        check_case(c, lambda req: oracle_lib.call(dict(req, indexed=True)))
