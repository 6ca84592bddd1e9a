"""The CPU oracle must reproduce every golden vector transcribed from the reference's own tests."""
import json
import os

import pytest

from tests import oracle_lib

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "reference_vectors.json")) as f:
    CASES = json.load(f)["cases"]


def _id(c):
    return "%s:%s:%s" % (c["op"], c.get("name", ""), c["test"][:60])


def check_case(c, call):
    req = {k: v for k, v in c.items() if k not in ("expect", "source", "test")}
    out = call(req)
    exp = c["expect"]
    if c["op"] == "priority":
        assert "error" not in out, out
        got = [[s["host"], s["score"]] for s in out["scores"]]
        assert got == exp["scores"], (c["source"], c["test"])
    elif c["op"] == "predicate":
        assert out.get("error", "") == exp.get("error", ""), (out, c["source"], c["test"])
        assert out["fits"] == exp["fits"], (out, c["source"], c["test"])
        if "reasons" in exp and exp["reasons"] is not None:
            assert sorted(out["reasons"]) == exp["reasons"], (c["source"], c["test"])
        if "devices" in exp:
            assert out["devices"] == exp["devices"], (c["source"], c["test"])
    elif c["op"] == "device_match":
        assert out["match"] == exp["match"], (out, c["source"], c["test"])
        assert bool(out["error"]) == exp["error"], (out, c["source"], c["test"])
    elif c["op"] == "stream":
        got = [r["host"] for r in out["results"]]
        if "hosts" in exp:
            assert got == exp["hosts"], (c["source"], c["test"])
        if "host_in" in exp:
            for g, allowed in zip(got, exp["host_in"]):
                assert g in allowed, (g, allowed, c["source"], c["test"])
        if "status" in exp:
            assert [r["status"] for r in out["results"]] == exp["status"], (out, c["source"], c["test"])
        if "scores" in exp:
            assert [r["score"] for r in out["results"]] == exp["scores"], (out, c["source"], c["test"])
        if "scores_not" in exp:
            for r, s in zip(out["results"], exp["scores_not"]):
                assert r["status"] == 0 and r["score"] != s, (out, c["source"], c["test"])
        if "message_contains" in exp:
            for r, parts in zip(out["results"], exp["message_contains"]):
                for part in parts:
                    assert part in r["message"], (r, c["source"], c["test"])
    elif c["op"] == "node_info":
        for k, v in exp.items():
            assert out[k] == v, (k, out, c["source"], c["test"])
    elif c["op"] == "preempt":
        assert not out.get("error"), out
        if "candidates" in exp:  # selectNodesForPreemption: node -> set of victims (generic_scheduler_test.go:575-601)
            got = {n: sorted(v.split("/", 1)[1] for v in vs) for n, vs in out["candidates"]}
            assert got == {n: sorted(vs) for n, vs in exp["candidates"].items()}, (out, c["source"], c["test"])
        if "node_in" in exp:  # pickOneNodeForPreemption: any of the listed nodes
            assert out["node"] in exp["node_in"], (out, c["source"], c["test"])
        if "victims" in exp:  # Preempt: the pods to evict on the chosen node (generic_scheduler_test.go:1202-1216)
            assert sorted(v.split("/", 1)[1] for v in out["victims"]) == sorted(exp["victims"]), (out, c["source"], c["test"])
        if "potential" in exp:  # nodesWherePreemptionMightHelp
            assert sorted(out["potential"]) == sorted(exp["potential"]), (out, c["source"], c["test"])
    elif c["op"] == "helper":
        assert "error" not in out, out
        assert out["result"] == exp["result"], (out, c["source"], c["test"])
    elif c["op"] == "cache_ops":
        assert out["errors"] == exp["errors"], (out["errors"], c["source"], c["test"])
        assert out["nodes"] == exp["nodes"], (out["nodes"], c["source"], c["test"])
    else:
        raise AssertionError("unknown op " + c["op"])


@pytest.mark.parametrize("case", CASES, ids=[_id(c) for c in CASES])
def test_oracle_matches_reference_vector(case):
    check_case(case, oracle_lib.call)
