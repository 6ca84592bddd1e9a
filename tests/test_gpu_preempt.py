"""ksb_preempt (genericScheduler.Preempt, core/generic_scheduler.go:199-256) against the oracle's restatement: the
reference's own preemption vectors (TestSelectNodesForPreemption, TestPickOneNodeForPreemption) and seeded clusters with
dozens of equal-priority pods per node, where the victims depend on the order go1.9's sort.Sort leaves them in."""
import ctypes
import json
import os

import pytest

from kubernetes_b200 import lib, wire, workloads
from tests import oracle_lib

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


class PreemptResult(ctypes.Structure):
    _fields_ = [("status", ctypes.c_int32), ("node_index", ctypes.c_int32), ("num_victims", ctypes.c_int32),
                ("num_potential", ctypes.c_int32), ("num_candidates", ctypes.c_int32)]


def _so():
    so = lib.engine()
    so.ksb_preempt.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(PreemptResult)]
    so.ksb_preempt_victim.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_size_t)]
    so.ksb_node_name.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_size_t)]
    so.ksb_selector_owner_upsert.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_char_p, ctypes.c_size_t]
    return so


def _create(so):
    cfg = lib.KsbConfig()
    so.ksb_config_default(ctypes.byref(cfg))
    err = ctypes.create_string_buffer(512)
    e = so.ksb_create(ctypes.byref(cfg), err, 512)
    assert e, err.value
    return e


def _preempt(so, e, pod):
    b = wire.pod_record(pod)
    r = PreemptResult()
    rc = so.ksb_preempt(e, b, len(b), ctypes.byref(r))
    if rc != 0:
        return {"error": so.ksb_last_error(e).decode()}
    out = {"status": r.status, "potential": r.num_potential, "candidates": r.num_candidates, "node": "", "victims": []}
    if r.status == 0:
        nm, nl = ctypes.c_char_p(), ctypes.c_size_t()
        assert so.ksb_node_name(e, r.node_index, ctypes.byref(nm), ctypes.byref(nl)) == 0
        out["node"] = nm.value[:nl.value].decode()
        for i in range(r.num_victims):
            k, kl = ctypes.c_char_p(), ctypes.c_size_t()
            assert so.ksb_preempt_victim(e, i, ctypes.byref(k), ctypes.byref(kl)) == 0
            out["victims"].append(k.value[:kl.value].decode())
    return out


def _load(so, e, cluster):
    for n in cluster["nodes"]:
        a = dict(n.get("allocatable", {}))
        a.setdefault("pods", 110)
        b = wire.node_record(dict(n, allocatable=a))
        assert so.ksb_node_upsert(e, b, len(b)) == 0, so.ksb_last_error(e)
    for p in cluster.get("existingPods", []):
        b = wire.pod_record(p)
        assert so.ksb_pod_add(e, b, len(b)) == 0, so.ksb_last_error(e)


def _check(so, cluster, pod, label):
    want = oracle_lib.call({"op": "preempt", "cluster": cluster, "pod": pod})
    e = _create(so)
    try:
        _load(so, e, cluster)
        got = _preempt(so, e, pod)
    finally:
        so.ksb_destroy(e)
    if want.get("error"):
        assert got.get("error") == want["error"], (label, got, want)
        return got, want
    assert "error" not in got, (label, got)
    assert got["potential"] == len(want["potential"]), (label, got, want)
    assert got["candidates"] == len(want["candidates"]), (label, got, want)
    assert got["node"] == want["node"], (label, got, want)
    assert got["victims"] == want["victims"], (label, got, want)  # same pods in the same (decreasing priority) order
    assert got["status"] == (0 if want["node"] else (2 if want["clear_nomination"] else 1)), (label, got, want)
    return got, want


def test_reference_preemption_vectors_on_gpu():
    with open(os.path.join(HERE, "golden", "reference_vectors.json")) as f:
        cases = [c for c in json.load(f)["cases"] if c["op"] == "preempt" and "failed" not in c]
    so = _so()
    assert len(cases) == 12
    for c in cases:
        cl = dict(c["cluster"])
        cl["nodes"] = [dict(n, allocatable=dict(n["allocatable"], pods=n["allocatable"].get("pods", 110))) for n in cl["nodes"]]
        got, want = _check(so, cl, c["pod"], c["test"])
        exp = c["expect"]
        if "node_in" in exp:  # pickOneNodeForPreemption: any of the listed nodes is right
            assert got["node"] in exp["node_in"], (c["test"], got)
        if "candidates" in exp:  # selectNodesForPreemption: node -> victims
            assert got["candidates"] == len(exp["candidates"]), (c["test"], got)
            if got["node"]:
                assert sorted(v.split("/", 1)[1] for v in got["victims"]) == sorted(exp["candidates"][got["node"]]), (c["test"], got)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_preemption_matches_oracle_on_crowded_nodes(seed):
    """40 nodes filled by the engine itself with pods of five priorities (about 21 per node, so the victims of a node go
    through the quicksort branch of the restated sort.Sort), then preemptors of several sizes and priorities."""
    so = _so()
    rng = workloads.SplitMix64(1000 + seed)
    cluster = workloads.generate(1, 40, 1800, seed=77 + seed)
    prios = [-100, 0, 100, 100, 500, 1000]
    for p in cluster["pods"]:
        p["priority"] = rng.pick(prios)
    e = _create(so)
    try:
        _load(so, e, {"nodes": cluster["nodes"]})
        recs = [wire.pod_record(p) for p in cluster["pods"]]
        offs = [0]
        for r in recs:
            offs.append(offs[-1] + len(r))
        out = (lib.KsbResult * len(recs))()
        assert so.ksb_schedule_batch(e, b"".join(recs), (ctypes.c_uint64 * (len(recs) + 1))(*offs), len(recs), out) == 0
        names = sorted(n["name"] for n in cluster["nodes"])
        placed = [dict(p, nodeName=names[out[i].node_index]) for i, p in enumerate(cluster["pods"]) if out[i].status == 0]
        assert len(placed) > 600  # the nodes are full: about 21 pods each
        state = {"nodes": cluster["nodes"], "existingPods": placed}
        checked = picked = 0
        for k in range(12):
            pod = {"name": "preemptor-%d" % k, "namespace": "default", "priority": rng.pick([50, 200, 700, 5000]),
                   "containers": [{"requests": {"cpu": 500 * (2 + rng.below(20)), "memory": (2 + rng.below(30)) << 30}}]}
            want = oracle_lib.call({"op": "preempt", "cluster": state, "pod": pod})
            assert not want.get("error"), want
            got = _preempt(so, e, pod)
            assert "error" not in got, got
            assert (got["potential"], got["candidates"], got["node"], got["victims"]) == (
                len(want["potential"]), len(want["candidates"]), want["node"], want["victims"]), (k, pod, got, want)
            checked += 1
            picked += 1 if got["node"] else 0
        assert checked == 12 and picked >= 4, (checked, picked)
    finally:
        so.ksb_destroy(e)


def test_preemption_out_of_scope_is_refused():
    so = _so()
    cluster = {"nodes": [{"name": "n1", "allocatable": {"cpu": 1000, "memory": 1 << 30, "pods": 10}}], "existingPods": []}
    pod = {"name": "p", "namespace": "default", "priority": 10,
           "containers": [{"requests": {"cpu": 100}, "ports": [{"hostPort": 8080, "protocol": "TCP"}]}]}
    got, want = _check(so, cluster, pod, "host ports")
    assert "host ports" in got["error"]
