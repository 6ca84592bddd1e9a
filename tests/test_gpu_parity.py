"""Parity proper: the CUDA engine, driven through the C-ABI, must place every pod exactly where the
CPU oracle (the restatement of the reference's Go code) places it, on the same seeded inputs.
Bar: bit-exact (hosts, device ids, filtered counts, scores, FitError text)."""
import json
import os

import pytest

from kubernetes_b200 import lib, workloads
from tests import oracle_lib

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def compare(cluster, label, gpu_opts=None, oracle_opts=None):
    gpu_opts = dict(gpu_opts or {})
    want = oracle_lib.call(dict({"op": "stream", "cluster": cluster}, **(oracle_opts or {})))
    assert "error" not in want, want
    sess = lib.Session(cluster=cluster)
    got = sess.run(**gpu_opts)
    sess.close()
    assert len(got["results"]) == len(want["results"])
    for i, (g, w) in enumerate(zip(got["results"], want["results"])):
        ctx = "%s pod %d" % (label, i)
        assert g["status"] == w["status"], (ctx, g, w)
        assert g["host"] == w["host"], (ctx, g, w)
        assert g["filtered"] == w["filtered"], (ctx, g, w)
        if w["status"] == 0:
            assert g["score"] == w["score"], (ctx, g, w)
            assert g["devices"] == w["devices"], (ctx, g, w)
        elif w["status"] == 2:
            assert g["message"] == w["message"], (ctx, g, w)
    assert got["lastNodeIndex"] == want["lastNodeIndex"], label
    return got, want


@pytest.mark.parametrize("cfg,nodes,pods", [(1, 100, 1000), (2, 150, 2000), (3, 400, 2500), (5, 150, 1200)])
def test_batch_matches_oracle(cfg, nodes, pods):
    cluster = workloads.generate(cfg, nodes, pods)
    compare(cluster, "cfg%d" % cfg)


@pytest.mark.parametrize("cfg,nodes,pods", [(1, 60, 300), (2, 40, 400), (3, 150, 500), (5, 80, 400)])
def test_single_pod_path_matches_oracle(cfg, nodes, pods):
    """ksb_schedule followed by ksb_pod_add (the reference's Schedule + AssumePod sequence)."""
    cluster = workloads.generate(cfg, nodes, pods)
    compare(cluster, "cfg%d single" % cfg, gpu_opts={"batch": 0})


def test_small_batches_match_one_batch():
    cluster = workloads.generate(3, 300, 1500)
    sess = lib.Session(cluster=cluster)
    a = sess.run(results=False)
    b = sess.run(results=False, batch=37)
    sess.close()
    assert a["placement_hash"] == b["placement_hash"]


@pytest.mark.parametrize("cfg,nodes,pods", [(1, 64, 800), (2, 60, 900), (3, 700, 3000), (5, 120, 900)])
def test_group_sizes_agree(cfg, nodes, pods):
    """Speculative groups of any size must give the placements of one-pod groups (live evaluation)."""
    cluster = workloads.generate(cfg, nodes, pods)
    sess = lib.Session(cluster=cluster)
    base = sess.run(results=False, group_size=1)
    seen = {}
    for g in (2, 5, 16, 32):
        out = sess.run(results=False, group_size=g)
        seen[g] = (out["placement_hash"], out["lastNodeIndex"])
        assert out["groups"] <= base["groups"]
    sess.close()
    assert all(v == (base["placement_hash"], base["lastNodeIndex"]) for v in seen.values()), (seen, base["placement_hash"])


@pytest.mark.parametrize("cfg,nodes,pods", [(1, 100, 1000), (2, 150, 2000), (3, 400, 2500), (5, 150, 1200)])
def test_pipelined_batch_matches_oracle(cfg, nodes, pods):
    """Commit of group k-1 overlapped with the evaluation of group k (forced on; automatic only from 16,384 rows)."""
    cluster = workloads.generate(cfg, nodes, pods)
    compare(cluster, "cfg%d pipelined" % cfg, gpu_opts={"pipeline": 1})


@pytest.mark.parametrize("cfg,nodes,pods", [(1, 64, 800), (2, 60, 900), (3, 700, 3000), (5, 120, 900)])
def test_pipelined_group_sizes_agree(cfg, nodes, pods):
    cluster = workloads.generate(cfg, nodes, pods)
    sess = lib.Session(cluster=cluster)
    base = sess.run(results=False, group_size=1, pipeline=-1)
    seen = {}
    for g in (2, 3, 7, 16):
        out = sess.run(results=False, group_size=g, pipeline=1)
        seen[g] = (out["placement_hash"], out["lastNodeIndex"])
    sess.close()
    assert all(v == (base["placement_hash"], base["lastNodeIndex"]) for v in seen.values()), (seen, base["placement_hash"])


@pytest.mark.parametrize("cfg,nodes,pods", [(3, 700, 3000), (5, 120, 900)])
def test_pipelined_is_repeatable(cfg, nodes, pods):
    """The pipelined schedule has blocks in different iterations at the same time; its published group sizes must not
    be overwritten early (iterations that evaluate nothing are the delicate case: cfg5 has many)."""
    cluster = workloads.generate(cfg, nodes, pods)
    sess = lib.Session(cluster=cluster)
    base = sess.run(results=False, pipeline=-1)
    for _ in range(6):
        for g in (4, 16):
            out = sess.run(results=False, pipeline=1, group_size=g)
            assert (out["placement_hash"], out["lastNodeIndex"]) == (base["placement_hash"], base["lastNodeIndex"])
    sess.close()


def test_pipelined_stress_many_runs():
    """Sixty pipelined runs at two group sizes: a data race between the deciding warps, the writer warp and the
    evaluating blocks shows up as a placement that changes from run to run (it did, once, during development)."""
    cluster = workloads.generate(3, 700, 3000)
    sess = lib.Session(cluster=cluster)
    base = sess.run(results=False, pipeline=-1, group_size=32)
    bad = 0
    for _ in range(60):
        for g in (4, 16):
            out = sess.run(results=False, pipeline=1, group_size=g)
            bad += (out["placement_hash"], out["lastNodeIndex"]) != (base["placement_hash"], base["lastNodeIndex"])
    sess.close()
    assert bad == 0


def test_chunked_batch_matches_single_launches():
    """Batches of 8,192 pods or more go to the device in chunks overlapped with host work; the placements must be
    those of one launch (the replay runs the whole batch in a single launch and is compared pod by pod) and of
    small separate batches."""
    cluster = workloads.generate(3, 1500, 26000)
    sess = lib.Session(cluster=cluster)
    a = sess.run(results=False, replay_iters=1)
    b = sess.run(results=False, batch=6000)
    sess.close()
    assert a["placement_hash"] == b["placement_hash"]
    assert a["lastNodeIndex"] == b["lastNodeIndex"]
    assert a["fit_errors"] == b["fit_errors"]


def test_chunked_batch_matches_oracle_with_fit_errors():
    """A chunked batch on a cluster that fills up: FitError texts of pods in later chunks come from reason histograms
    that are collected per launch and renumbered."""
    cluster = workloads.generate(1, 120, 14000)
    got, want = compare(cluster, "cfg1 chunked", oracle_opts={"workers": 8})
    assert sum(1 for r in want["results"] if r["status"] == 2) > 1000


def test_pipelined_tight_cluster():
    cluster = workloads.generate(1, 12, 700)
    for n in cluster["nodes"]:
        n["allocatable"]["pods"] = 40
    got, want = compare(cluster, "tight cfg1 pipelined", gpu_opts={"pipeline": 1})
    assert any(r["status"] == 2 for r in want["results"])


def test_full_node_and_tight_cluster_forces_restarts():
    """A small, nearly full cluster: rows turn unfit inside groups, so speculation is dropped often."""
    cluster = workloads.generate(1, 12, 700)
    for n in cluster["nodes"]:
        n["allocatable"]["pods"] = 40
    got, want = compare(cluster, "tight cfg1")
    assert any(r["status"] == 2 for r in want["results"])
    assert got["group_restarts"] > 0


def test_unpadded_names_and_heterogeneous_devices():
    c1 = workloads.generate(1, 120, 600, padded=False)  # "node-10" < "node-9" bytewise
    compare(c1, "cfg1 unpadded")
    c2 = workloads.generate(2, 120, 1500, hetero_pct=10)
    compare(c2, "cfg2 hetero")


def test_golden_stream_cases_on_gpu():
    with open(os.path.join(HERE, "golden", "reference_vectors.json")) as f:
        cases = [c for c in json.load(f)["cases"] if c["op"] == "stream" and "extender" not in c]  # (those: test_gpu_extender.py)
    assert cases
    for c in cases:
        sess = lib.Session(cluster=c["cluster"])
        out = sess.run(weights=c.get("weights"))
        sess.close()
        exp = c["expect"]
        if "scores" in exp:
            assert [r["score"] for r in out["results"]] == exp["scores"], c["test"]
        if "host_in" in exp:
            for r, allowed in zip(out["results"], exp["host_in"]):
                assert r["host"] in allowed
        if "status" in exp:
            assert [r["status"] for r in out["results"]] == exp["status"], c["test"]
        if "scores_not" in exp:
            for r, s in zip(out["results"], exp["scores_not"]):
                assert r["status"] == 0 and r["score"] != s, (c["test"], r)
        if "message_contains" in exp:
            for r, parts in zip(out["results"], exp["message_contains"]):
                for part in parts:
                    assert part in r["message"], (c["test"], r)


def _as_stream(case):
    """Run a golden predicate case through the whole engine: one node listed, one pod."""
    cl = dict(case["cluster"])
    cl["nodes"] = [dict(n) for n in cl["nodes"]]
    for n in cl["nodes"]:
        a = dict(n.get("allocatable", {}))
        a.setdefault("pods", 110)
        n["allocatable"] = a
    cl["pods"] = [case["pod"]]
    return cl


def test_golden_predicates_on_gpu():
    """Every transcribed predicate vector, pushed through Schedule on a cluster that lists only the
    node under test (other nodes only carry existing pods): fit <=> the pod is placed there."""
    with open(os.path.join(HERE, "golden", "reference_vectors.json")) as f:
        cases = [c for c in json.load(f)["cases"] if c["op"] == "predicate"]
    checked = 0
    for c in cases:
        if c["expect"].get("error"):
            continue
        cl = _as_stream(c)
        policy = c.get("policy")  # CheckNodeLabelPresence takes its labels from a Policy
        want = oracle_lib.call({"op": "stream", "cluster": cl, "policy": policy})
        sess = lib.Session(cluster=cl)
        got = sess.run(policy=policy) if policy else sess.run()
        sess.close()
        g, w = got["results"][0], want["results"][0]
        assert (g["status"], g["host"], g["devices"]) == (w["status"], w["host"], w["devices"]), (c["source"], c["test"], g, w)
        if w["status"] == 2:
            assert g["message"] == w["message"], (c["source"], c["test"])
        checked += 1
    assert checked > 100


def test_golden_priorities_on_gpu():
    """Every transcribed priority vector: schedule the pod with only that priority weighted and compare the WHOLE
    HostPriorityList (ksb_scores) with the golden list, as the reference's tests do with reflect.DeepEqual
    (priorities/test_util.go:45-61). Cases in which a predicate removes a node are counted, not silently skipped:
    the golden lists are computed without predicates, so only the nodes that stayed are compared there."""
    names = {"LeastRequestedPriority", "BalancedResourceAllocation", "SelectorSpreadPriority", "NodeAffinityPriority",
             "TaintTolerationPriority", "InterPodAffinityPriority", "NodePreferAvoidPodsPriority", "MostRequestedPriority",
             "ResourceLimitsPriority", "ImageLocalityPriority"}
    with open(os.path.join(HERE, "golden", "reference_vectors.json")) as f:
        cases = [c for c in json.load(f)["cases"] if c["op"] == "priority"]
    whole = partial = skipped = image_whole = 0
    for c in cases:
        cl = _as_stream(c)
        for n in cl["nodes"]:  # priorities are tested without predicates in the reference: make every node fit
            n["allocatable"]["cpu"] = max(n["allocatable"].get("cpu", 0), 0)
        weights = {k: (1 if k == c["name"] else 0) for k in names}
        scores = dict((h, s) for h, s in c["expect"]["scores"])
        best = max(scores.values())
        winners = {h for h, s in scores.items() if s == best}
        policy = None
        if c["name"] == "NodeLabelPriority":  # an instance with arguments (node_label.go:29-44), not a weight by name
            policy = {"nodeLabelPriorities": [dict(c["args"], weight=1)]}
        sess = lib.Session(cluster=cl)
        out = sess.run(batch=0, scores=True, weights=weights, hardPodAffinityWeight=c.get("hardPodAffinityWeight", 1),
                       **({"policy": policy} if policy else {}))
        sess.close()
        r = out["results"][0]
        if r["status"] != 0 or r["filtered"] < 2:
            skipped += 1  # fewer than two nodes fit: the reference does not prioritize at all
            continue
        got = r["scores"]
        if r["filtered"] == len(scores):
            assert got == scores, (c["source"], c["test"], got, scores)
            assert r["host"] in winners and r["score"] == best, (c["source"], c["test"], r, scores)
            whole += 1
            image_whole += c["name"] == "ImageLocalityPriority"
        else:
            # a predicate removed nodes; priorities whose Reduce normalises over the filtered list may legitimately
            # differ from the golden list then, the Map-only ones may not
            if c["name"] in ("LeastRequestedPriority", "BalancedResourceAllocation", "NodePreferAvoidPodsPriority",
                             "MostRequestedPriority", "ResourceLimitsPriority", "NodeLabelPriority", "ImageLocalityPriority"):
                assert all(got[h] == scores[h] for h in got), (c["source"], c["test"], got, scores)
            partial += 1
    assert whole + partial + skipped == len(cases)
    print("priority vectors: %d whole lists, %d partial, %d not prioritized" % (whole, partial, skipped))
    assert whole >= 85 and image_whole == 3, (whole, partial, skipped, image_whole)


def _many_sets_cluster(n_nodes=300, n_pods=1200, n_hard=150, n_soft=140, n_avoid=130):
    """cfg3 with more than 64 distinct NoSchedule taints, PreferNoSchedule taints and preferAvoidPods references."""
    cluster = workloads.generate(3, n_nodes, n_pods)
    for i, node in enumerate(cluster["nodes"]):
        taints = list(node.get("taints") or [])
        if i % 3 == 0:
            taints.append({"key": "hard-%d" % ((i // 3) % n_hard), "value": "x", "effect": "NoSchedule"})  # 100 distinct at 300 nodes
        if i % 2 == 0:
            taints.append({"key": "soft-%d" % (i % n_soft), "value": "y", "effect": "PreferNoSchedule"})
            taints.append({"key": "soft-%d" % ((7 * i + 3) % n_soft), "value": "y", "effect": "PreferNoSchedule"})
        node["taints"] = taints
        node["avoid"] = [{"kind": "ReplicationController", "uid": "rc-%d" % ((i + j) % n_avoid)} for j in range(3)]
    for i, pod in enumerate(cluster["pods"]):
        tols = list(pod.get("tolerations") or [])
        for j in range(40):  # tolerate a spread of the hard taints, so that many nodes stay feasible
            tols.append({"key": "hard-%d" % ((i + 4 * j) % n_hard), "operator": "Exists", "value": "", "effect": "NoSchedule"})
        for j in range(5):
            tols.append({"key": "soft-%d" % ((3 * i + j) % n_soft), "operator": "Equal", "value": "y", "effect": "PreferNoSchedule"})
        pod["tolerations"] = tols
        pod["controller"] = {"kind": "ReplicationController" if i % 2 else "ReplicaSet", "uid": "rc-%d" % (i % n_avoid)}
    return cluster


def test_more_than_64_taints_and_avoid_references():
    """The taint / preferAvoid sets grow to several 64-bit words (first 64 entries in the header word, the rest behind
    the pod's pool block): batch and one-pod paths against the oracle."""
    cluster = _many_sets_cluster()
    compare(cluster, "many taints")
    compare(_many_sets_cluster(120, 300), "many taints single", gpu_opts={"batch": 0})


def test_policy_plugins_match_oracle():
    """A Policy that enables MostRequestedPriority, ResourceLimitsPriority, two NodeLabelPriority instances and the
    CheckNodeLabelPresence predicate next to the DefaultProvider set: whole streams (batch and one pod per call)
    against the oracle."""
    cluster = workloads.generate(3, 300, 1500)
    for i, p in enumerate(cluster["pods"]):  # limits on some pods, so that ResourceLimitsPriority has something to score
        if i % 3 == 0:
            p["containers"][0]["limits"] = {"cpu": 500 * (1 + i % 80), "memory": (1 + i % 5) << 30}
    opts = {"weights": {"LeastRequestedPriority": 1, "BalancedResourceAllocation": 1, "SelectorSpreadPriority": 1,
                        "NodeAffinityPriority": 1, "TaintTolerationPriority": 1, "InterPodAffinityPriority": 1,
                        "NodePreferAvoidPodsPriority": 10000, "MostRequestedPriority": 2, "ResourceLimitsPriority": 3},
            "policy": {"nodeLabelPriorities": [{"label": "lk03", "presence": True, "weight": 2},
                                               {"label": "lk07", "presence": False, "weight": 1}],
                       "labelsPresence": {"labels": ["lk11"], "presence": False}}}
    compare(cluster, "policy", gpu_opts=opts, oracle_opts=opts)
    small = workloads.generate(3, 120, 300)
    compare(small, "policy single", gpu_opts=dict(opts, batch=0), oracle_opts=opts)
    # the two legacy names: ServiceSpreadingPriority scores as SelectorSpread does once a priorityMetadata exists
    # (selector_spreading.go:71-76), EqualPriority is 1 on every node (generic_scheduler.go:643-652)
    legacy = {"weights": dict(opts["weights"], ServiceSpreadingPriority=2, EqualPriority=3, MostRequestedPriority=0,
                              ResourceLimitsPriority=0)}
    got, want = compare(workloads.generate(3, 200, 800), "legacy names", gpu_opts=legacy, oracle_opts=legacy)
    plain = oracle_lib.call({"op": "stream", "cluster": workloads.generate(3, 200, 800)})
    assert any(w["score"] != p["score"] for w, p in zip(want["results"], plain["results"]))
    only_equal = {"weights": {k: 0 for k in opts["weights"]}}
    only_equal["weights"]["EqualPriority"] = 1  # a Policy without priorities: every node scores 1, round-robin over all that fit
    compare(workloads.generate(1, 50, 200), "equal only", gpu_opts=dict(only_equal, batch=0), oracle_opts=only_equal)
    compare(workloads.generate(1, 50, 200), "equal only batch", gpu_opts=only_equal, oracle_opts=only_equal)


def test_label_universe_grows_past_256_pairs():
    """More distinct (key, value) pairs in selectors than four 64-bit label words hold: the engine doubles the label
    columns (GrowLabels) and goes on; the general label-word path of the kernel against the oracle."""
    cluster = workloads.generate(3, 240, 900)
    for i, n in enumerate(cluster["nodes"]):
        for k in range(6):
            n["labels"]["grp%02d" % ((i + 7 * k) % 45)] = "val%d" % ((i // 3 + k) % 9)  # 405 possible pairs
    for i, p in enumerate(cluster["pods"]):
        p["nodeSelector"] = {}
        p["affinity"] = {"node": {"required": [{"matchExpressions": [
            {"key": "grp%02d" % ((3 * i + j) % 45), "op": "In", "values": ["val%d" % ((i + j + v) % 9) for v in range(5)]}
            for j in range(2)]}, {"matchExpressions": [{"key": "grp%02d" % ((5 * i) % 45), "op": "Exists", "values": []}]}]}}
    compare(cluster, "wide labels")
    compare(cluster, "wide labels single", gpu_opts={"batch": 0, "limit": 300}, oracle_opts={"limit": 300})


def test_label_dictionary_starts_afresh_when_full():
    """Selectors that mention more than 1,024 distinct pairs over time (never in one call): at 16 label words the
    dictionary is rebuilt from what the current call needs instead of failing; results stay those of the oracle."""
    cluster = workloads.generate(1, 120, 0)
    for i, n in enumerate(cluster["nodes"]):
        for k in range(40):
            n["labels"]["k%03d" % ((i * 7 + k * 13) % 300)] = "v%d" % ((i + k) % 8)
    pods = []
    for b in range(5):  # five batches, each mentioning about 480 pairs of its own key range
        for i in range(60):
            p = {"name": "p-%d-%d" % (b, i), "namespace": "default", "containers": [{"requests": {"cpu": 100, "memory": 1 << 27}}]}
            p["affinity"] = {"node": {"required": [{"matchExpressions": [
                {"key": "k%03d" % (60 * b + (i + j) % 60), "op": "In", "values": ["v%d" % ((i + j + v) % 8) for v in range(4)]}
                for j in range(2)]}]}}
            pods.append(p)
    cluster["pods"] = pods
    want = oracle_lib.call({"op": "stream", "cluster": cluster})
    sess = lib.Session(cluster=cluster)
    got = sess.run(batch=60)
    sess.close()
    for i, (g, w) in enumerate(zip(got["results"], want["results"])):
        assert (g["status"], g["host"], g["filtered"]) == (w["status"], w["host"], w["filtered"]), (i, g, w)
    assert sum(1 for r in want["results"] if r["status"] == 0) > 100
