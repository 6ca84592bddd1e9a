#!/usr/bin/env python3
"""Transcription of the reference's own table-driven tests for the scheduler hot
path into JSON golden vectors (tests/golden/reference_vectors.json).

Every case cites the reference test it was transcribed from (file:line under
/root/reference/plugin/pkg/scheduler). Expected values are the ones the Go tests
assert with reflect.DeepEqual / require.*; nothing here is computed by our code.
Run:  python tests/golden/build_fixtures.py
"""
import json
import os

CASES = []
_uid = [0]

ZONE = "failure-domain.beta.kubernetes.io/zone"
REGION = "failure-domain.beta.kubernetes.io/region"
HOSTNAME = "kubernetes.io/hostname"


def _name(prefix):
    _uid[0] += 1
    return "%s-%d" % (prefix, _uid[0])


def node(name, cpu=None, mem=None, labels=None, pods=None, alloc=None, **kw):
    a = {}
    if cpu is not None:
        a["cpu"] = cpu
    if mem is not None:
        a["memory"] = mem
    if pods is not None:
        a["pods"] = pods
    if alloc:
        a.update(alloc)
    n = {"name": name, "allocatable": a}
    if labels is not None:
        n["labels"] = labels
    n.update(kw)
    return n


def ctr(cpu=None, mem=None, requests=None, limits=None, ports=None):
    r = {}
    if cpu is not None:
        r["cpu"] = cpu
    if mem is not None:
        r["memory"] = mem
    if requests:
        r.update(requests)
    c = {"requests": r}
    if limits:
        c["limits"] = limits
    if ports:
        c["ports"] = ports
    return c


def pod(node_name=None, labels=None, ns="", containers=None, name=None, **kw):
    p = {"name": name or _name("pod"), "namespace": ns}
    if labels is not None:
        p["labels"] = labels
    if node_name:
        p["nodeName"] = node_name
    if containers is not None:
        p["containers"] = containers
    p.update(kw)
    return p


def expr(key, op, *values):
    return {"key": key, "op": op, "values": list(values)}


def priority(name, src, test, pod_, nodes, pods=None, expect=None, **cluster_kw):
    cl = {"nodes": nodes, "existingPods": pods or []}
    cl.update(cluster_kw)
    CASES.append({"op": "priority", "name": name, "source": src, "test": test, "cluster": cl, "pod": pod_,
                  "expect": {"scores": [[h, s] for h, s in expect]}})


def predicate(name, src, test, pod_, node_, node_pods=None, fits=True, reasons=None, other_nodes=None,
              other_pods=None, **cluster_kw):
    nodes = [node_] + (other_nodes or [])
    pods = []
    for p in (node_pods or []):
        q = dict(p)
        q["nodeName"] = node_["name"]
        pods.append(q)
    pods += other_pods or []
    cl = {"nodes": nodes, "existingPods": pods}
    cl.update(cluster_kw)
    CASES.append({"op": "predicate", "name": name, "source": src, "test": test, "cluster": cl, "pod": pod_,
                  "node": node_["name"], "expect": {"fits": fits, "reasons": sorted(reasons or [])}})


# --------------------------------------------------------------------------- LeastRequested / Balanced
# algorithm/priorities/least_requested_test.go:30-268, balanced_resource_allocation_test.go:30-264

def _lr_specs():
    labels1 = {"foo": "bar", "baz": "blah"}
    labels2 = {"bar": "foo", "baz": "blah"}
    cpu_only = [ctr(1000, 0), ctr(2000, 0)]
    cpu_mem = [ctr(1000, 2000), ctr(2000, 3000)]
    return labels1, labels2, cpu_only, cpu_mem


def build_resource_priorities():
    l1, l2, cpu_only, cpu_mem = _lr_specs()
    no_res = pod(containers=[])
    P = lambda n, c=None, l=None: pod(n, labels=l, containers=c if c is not None else [])
    tables = [
        ("nothing scheduled, nothing requested", no_res, [("machine1", 4000, 10000), ("machine2", 4000, 10000)], [],
         (10, 10), (10, 10)),
        ("nothing scheduled, resources requested, differently sized machines", pod(containers=cpu_mem),
         [("machine1", 4000, 10000), ("machine2", 6000, 10000)], [], (3, 5), (7, 10)),
        ("no resources requested, pods scheduled", no_res, [("machine1", 4000, 10000), ("machine2", 4000, 10000)],
         [P("machine1", None, l2), P("machine1", None, l1), P("machine2", None, l1), P("machine2", None, l1)],
         (10, 10), (10, 10)),
        ("no resources requested, pods scheduled with resources", no_res,
         [("machine1", 10000, 20000), ("machine2", 10000, 20000)],
         [P("machine1", cpu_only, l2), P("machine1", cpu_only, l1), P("machine2", cpu_only, l1),
          P("machine2", cpu_mem, l1)], (7, 5), (4, 6)),
        ("resources requested, pods scheduled with resources", pod(containers=cpu_mem),
         [("machine1", 10000, 20000), ("machine2", 10000, 20000)], [P("machine1", cpu_only), P("machine2", cpu_mem)],
         (5, 4), (6, 9)),
        ("resources requested, pods scheduled with resources, differently sized machines", pod(containers=cpu_mem),
         [("machine1", 10000, 20000), ("machine2", 10000, 50000)], [P("machine1", cpu_only), P("machine2", cpu_mem)],
         (5, 6), (6, 6)),
        ("requested resources exceed node capacity", pod(containers=cpu_only),
         [("machine1", 4000, 10000), ("machine2", 4000, 10000)], [P("machine1", cpu_only), P("machine2", cpu_mem)],
         (5, 2), (0, 0)),
        ("zero node resources, pods scheduled with resources", no_res, [("machine1", 0, 0), ("machine2", 0, 0)],
         [P("machine1", cpu_only), P("machine2", cpu_mem)], (0, 0), (0, 0)),
    ]
    for test, p, nodes, pods, lr, bra in tables:
        ns = [node(n, c, m) for n, c, m in nodes]
        priority("LeastRequestedPriority", "algorithm/priorities/least_requested_test.go:94-255", test, p, ns, pods,
                 list(zip([n[0] for n in nodes], lr)))
        priority("BalancedResourceAllocation", "algorithm/priorities/balanced_resource_allocation_test.go:99-252", test,
                 p, ns, pods, list(zip([n[0] for n in nodes], bra)))


# --------------------------------------------------------------------------- SelectorSpread
# algorithm/priorities/selector_spreading_test.go:43-365 and :375-603

def cref(kind, uid="abc123"):
    return {"kind": kind, "uid": uid}


def build_selector_spread():
    src = "algorithm/priorities/selector_spreading_test.go:43-365"
    l1 = {"foo": "bar", "baz": "blah"}
    l2 = {"bar": "foo", "baz": "blah"}
    nodes = lambda: [node("machine1"), node("machine2")]
    svc = lambda sel, ns="": {"namespace": ns, "selector": sel}
    ls = lambda m: {"namespace": "", "selector": {"matchLabels": m}}
    M1, M2 = "machine1", "machine2"
    T = [
        ("nothing scheduled", pod(), [], {}, (10, 10)),
        ("no services", pod(labels=l1), [pod(M1)], {}, (10, 10)),
        ("different services", pod(labels=l1), [pod(M1, l2)], {"services": [svc({"key": "value"})]}, (10, 10)),
        ("two pods, one service pod", pod(labels=l1), [pod(M1, l2), pod(M2, l1)], {"services": [svc(l1)]}, (10, 0)),
        ("five pods, one service pod in no namespace", pod(labels=l1),
         [pod(M1, l2), pod(M1, l1, ns="default"), pod(M1, l1, ns="ns1"), pod(M2, l1), pod(M2, l2)],
         {"services": [svc(l1)]}, (10, 0)),
        ("four pods, one service pod in default namespace", pod(labels=l1, ns="default"),
         [pod(M1, l1), pod(M1, l1, ns="ns1"), pod(M2, l1, ns="default"), pod(M2, l2)],
         {"services": [svc(l1, "default")]}, (10, 0)),
        ("five pods, one service pod in specific namespace", pod(labels=l1, ns="ns1"),
         [pod(M1, l1), pod(M1, l1, ns="default"), pod(M1, l1, ns="ns2"), pod(M2, l1, ns="ns1"), pod(M2, l2)],
         {"services": [svc(l1, "ns1")]}, (10, 0)),
        ("three pods, two service pods on different machines", pod(labels=l1), [pod(M1, l2), pod(M1, l1), pod(M2, l1)],
         {"services": [svc(l1)]}, (0, 0)),
        ("four pods, three service pods", pod(labels=l1), [pod(M1, l2), pod(M1, l1), pod(M2, l1), pod(M2, l1)],
         {"services": [svc(l1)]}, (5, 0)),
        ("service with partial pod label matches", pod(labels=l1), [pod(M1, l2), pod(M1, l1), pod(M2, l1)],
         {"services": [svc({"baz": "blah"})]}, (0, 5)),
        ("service with partial pod label matches with service and replication controller",
         pod(labels=l1, controller=cref("ReplicationController")),
         [pod(M1, l2), pod(M1, l1, controller=cref("ReplicationController")),
          pod(M2, l1, controller=cref("ReplicationController"))],
         {"services": [svc({"baz": "blah"})], "rcs": [svc({"foo": "bar"})]}, (0, 5)),
        ("service with partial pod label matches with service and replica set",
         pod(labels=l1, controller=cref("ReplicaSet")),
         [pod(M1, l2), pod(M1, l1, controller=cref("ReplicaSet")), pod(M2, l1, controller=cref("ReplicaSet"))],
         {"services": [svc({"baz": "blah"})], "rss": [ls({"foo": "bar"})]}, (0, 5)),
        ("service with partial pod label matches with service and stateful set",
         pod(labels=l1, controller=cref("StatefulSet")),
         [pod(M1, l2), pod(M1, l1, controller=cref("StatefulSet")), pod(M2, l1, controller=cref("StatefulSet"))],
         {"services": [svc({"baz": "blah"})], "sss": [ls({"foo": "bar"})]}, (0, 5)),
        ("disjoined service and replication controller should be treated equally",
         pod(labels={"foo": "bar", "bar": "foo"}, controller=cref("ReplicationController")),
         [pod(M1, l2), pod(M1, l1, controller=cref("ReplicationController")),
          pod(M2, l1, controller=cref("ReplicationController"))],
         {"services": [svc({"bar": "foo"})], "rcs": [svc({"foo": "bar"})]}, (0, 5)),
        ("disjoined service and replica set should be treated equally",
         pod(labels={"foo": "bar", "bar": "foo"}, controller=cref("ReplicaSet")),
         [pod(M1, l2), pod(M1, l1, controller=cref("ReplicaSet")), pod(M2, l1, controller=cref("ReplicaSet"))],
         {"services": [svc({"bar": "foo"})], "rss": [ls({"foo": "bar"})]}, (0, 5)),
        ("disjoined service and stateful set should be treated equally",  # (named "... replica set ..." a second time in the reference, :249-261)
         pod(labels={"foo": "bar", "bar": "foo"}, controller=cref("StatefulSet")),
         [pod(M1, l2), pod(M1, l1, controller=cref("StatefulSet")), pod(M2, l1, controller=cref("StatefulSet"))],
         {"services": [svc({"bar": "foo"})], "sss": [ls({"foo": "bar"})]}, (0, 5)),
        ("Replication controller with partial pod label matches", pod(labels=l1, controller=cref("ReplicationController")),
         [pod(M1, l2), pod(M1, l1, controller=cref("ReplicationController")),
          pod(M2, l1, controller=cref("ReplicationController"))], {"rcs": [svc({"foo": "bar"})]}, (0, 0)),
        ("Replica set with partial pod label matches", pod(labels=l1, controller=cref("ReplicaSet")),
         [pod(M1, l2), pod(M1, l1, controller=cref("ReplicaSet")), pod(M2, l1, controller=cref("ReplicaSet"))],
         {"rss": [ls({"foo": "bar"})]}, (0, 0)),
        ("StatefulSet with partial pod label matches", pod(labels=l1, controller=cref("StatefulSet")),
         [pod(M1, l2), pod(M1, l1, controller=cref("StatefulSet")), pod(M2, l1, controller=cref("StatefulSet"))],
         {"sss": [ls({"foo": "bar"})]}, (0, 0)),
        ("Another replication controller with partial pod label matches",
         pod(labels=l1, controller=cref("ReplicationController")),
         [pod(M1, l2, controller=cref("ReplicationController")), pod(M1, l1, controller=cref("ReplicationController")),
          pod(M2, l1, controller=cref("ReplicationController"))], {"rcs": [svc({"baz": "blah"})]}, (0, 5)),
        ("Another replication set with partial pod label matches", pod(labels=l1, controller=cref("ReplicaSet")),
         [pod(M1, l2, controller=cref("ReplicaSet")), pod(M1, l1, controller=cref("ReplicaSet")),
          pod(M2, l1, controller=cref("ReplicaSet"))], {"rss": [ls({"baz": "blah"})]}, (0, 5)),
        ("Another stateful set with partial pod label matches", pod(labels=l1, controller=cref("StatefulSet")),
         [pod(M1, l2, controller=cref("StatefulSet")), pod(M1, l1, controller=cref("StatefulSet")),
          pod(M2, l1, controller=cref("StatefulSet"))], {"sss": [ls({"baz": "blah"})]}, (0, 5)),
    ]
    for test, p, pods, kw, exp in T:
        priority("SelectorSpreadPriority", src, test, p, nodes(), pods, [(M1, exp[0]), (M2, exp[1])], **kw)

    # zone cases :375-603
    src = "algorithm/priorities/selector_spreading_test.go:375-603"
    l1 = {"label1": "l1", "baz": "blah"}
    l2 = {"label2": "l2", "baz": "blah"}
    names = ["machine1.zone1", "machine1.zone2", "machine2.zone2", "machine1.zone3", "machine2.zone3", "machine3.zone3"]
    m1z1, m1z2, m2z2, m1z3, m2z3, m3z3 = names
    znodes = lambda: [node(n, labels={ZONE: n.split(".")[1]}) for n in names]
    rc = cref("ReplicationController")
    Z = [
        ("nothing scheduled", pod(), [], {}, (10, 10, 10, 10, 10, 10)),
        ("no services", pod(labels=l1), [pod(m1z1)], {}, (10, 10, 10, 10, 10, 10)),
        ("different services", pod(labels=l1), [pod(m1z1, l2)], {"services": [svc({"key": "value"})]},
         (10, 10, 10, 10, 10, 10)),
        ("two pods, 0 matching", pod(labels=l1), [pod(m1z1, l2), pod(m1z2, l2)], {"services": [svc(l1)]},
         (10, 10, 10, 10, 10, 10)),
        ("two pods, 1 matching (in z2)", pod(labels=l1), [pod(m1z1, l2), pod(m1z2, l1)], {"services": [svc(l1)]},
         (10, 0, 3, 10, 10, 10)),
        ("five pods, 3 matching (z2=2, z3=1)", pod(labels=l1),
         [pod(m1z1, l2), pod(m1z2, l1), pod(m2z2, l1), pod(m1z3, l2), pod(m2z3, l1)], {"services": [svc(l1)]},
         (10, 0, 0, 6, 3, 6)),
        ("four pods, 3 matching (z1=1, z2=1, z3=1)", pod(labels=l1),
         [pod(m1z1, l1), pod(m1z2, l1), pod(m2z2, l2), pod(m1z3, l1)], {"services": [svc(l1)]}, (0, 0, 3, 0, 3, 3)),
        ("four pods, 3 matching (z1=1, z2=1, z3=1) [reordered]", pod(labels=l1),
         [pod(m1z1, l1), pod(m1z2, l1), pod(m1z3, l1), pod(m2z2, l2)], {"services": [svc(l1)]}, (0, 0, 3, 0, 3, 3)),
        ("Replication controller spreading (z1=0, z2=1, z3=2)", pod(labels=l1, controller=rc),
         [pod(m1z3, l1, controller=rc), pod(m1z2, l1, controller=rc), pod(m1z3, l1, controller=rc)],
         {"rcs": [svc(l1)]}, (10, 5, 6, 0, 3, 3)),
    ]
    for test, p, pods, kw, exp in Z:
        priority("SelectorSpreadPriority", src, test, p, znodes(), pods, list(zip(names, exp)), **kw)


# --------------------------------------------------------------------------- NodeAffinity priority
# algorithm/priorities/node_affinity_test.go:30-180

def build_node_affinity_priority():
    src = "algorithm/priorities/node_affinity_test.go:30-180"
    label1, label2, label3 = {"foo": "bar"}, {"key": "value"}, {"az": "az1"}
    label4 = {"abc": "az11", "def": "az22"}
    label5 = {"foo": "bar", "key": "value", "az": "az1"}
    aff1 = {"node": {"preferred": [{"weight": 2, "matchExpressions": [expr("foo", "In", "bar")]}]}}
    aff2 = {"node": {"preferred": [
        {"weight": 2, "matchExpressions": [expr("foo", "In", "bar")]},
        {"weight": 4, "matchExpressions": [expr("key", "In", "value")]},
        {"weight": 5, "matchExpressions": [expr("foo", "In", "bar"), expr("key", "In", "value"), expr("az", "In", "az1")]},
    ]}}
    N = lambda n, l: node(n, labels=l)
    priority("NodeAffinityPriority", src, "all machines are same priority as NodeAffinity is nil", pod(),
             [N("machine1", label1), N("machine2", label2), N("machine3", label3)], [],
             [("machine1", 0), ("machine2", 0), ("machine3", 0)])
    priority("NodeAffinityPriority", src, "no machine matches preferred scheduling requirements", pod(affinity=aff1),
             [N("machine1", label4), N("machine2", label2), N("machine3", label3)], [],
             [("machine1", 0), ("machine2", 0), ("machine3", 0)])
    priority("NodeAffinityPriority", src, "only machine1 matches the preferred scheduling requirements of pod",
             pod(affinity=aff1), [N("machine1", label1), N("machine2", label2), N("machine3", label3)], [],
             [("machine1", 10), ("machine2", 0), ("machine3", 0)])
    priority("NodeAffinityPriority", src, "all machines match with different priorities", pod(affinity=aff2),
             [N("machine1", label1), N("machine5", label5), N("machine2", label2)], [],
             [("machine1", 1), ("machine5", 10), ("machine2", 3)])


# --------------------------------------------------------------------------- TaintToleration priority
# algorithm/priorities/taint_toleration_test.go:50-230

def build_taint_priority():
    src = "algorithm/priorities/taint_toleration_test.go:50-230"
    T = lambda k, v, e: {"key": k, "value": v, "effect": e}
    Tol = lambda k, v, e: {"key": k, "operator": "Equal", "value": v, "effect": e}
    PNS, NS = "PreferNoSchedule", "NoSchedule"
    N = lambda n, taints: node(n, taints=taints)
    priority("TaintTolerationPriority", src, "tolerated taints score higher than intolerable",
             pod(tolerations=[Tol("foo", "bar", PNS)]),
             [N("nodeA", [T("foo", "bar", PNS)]), N("nodeB", [T("foo", "blah", PNS)])], [], [("nodeA", 10), ("nodeB", 0)])
    priority("TaintTolerationPriority", src, "all taints tolerated: same score regardless of count",
             pod(tolerations=[Tol("cpu-type", "arm64", PNS), Tol("disk-type", "ssd", PNS)]),
             [N("nodeA", []), N("nodeB", [T("cpu-type", "arm64", PNS)]),
              N("nodeC", [T("cpu-type", "arm64", PNS), T("disk-type", "ssd", PNS)])], [],
             [("nodeA", 10), ("nodeB", 10), ("nodeC", 10)])
    priority("TaintTolerationPriority", src, "the more intolerable taints a node has, the lower score it gets",
             pod(tolerations=[Tol("foo", "bar", PNS)]),
             [N("nodeA", []), N("nodeB", [T("cpu-type", "arm64", PNS)]),
              N("nodeC", [T("cpu-type", "arm64", PNS), T("disk-type", "ssd", PNS)])], [],
             [("nodeA", 10), ("nodeB", 5), ("nodeC", 0)])
    priority("TaintTolerationPriority", src, "only PreferNoSchedule taints and tolerations are checked",
             pod(tolerations=[Tol("cpu-type", "arm64", NS), Tol("disk-type", "ssd", NS)]),
             [N("nodeA", []), N("nodeB", [T("cpu-type", "arm64", NS)]),
              N("nodeC", [T("cpu-type", "arm64", PNS), T("disk-type", "ssd", PNS)])], [],
             [("nodeA", 10), ("nodeB", 10), ("nodeC", 0)])
    priority("TaintTolerationPriority", src, "no taints and tolerations: lands on node with no taints",
             pod(tolerations=[]), [N("nodeA", []), N("nodeB", [T("cpu-type", "arm64", PNS)])], [],
             [("nodeA", 10), ("nodeB", 0)])


BUILDERS = [build_resource_priorities, build_selector_spread, build_node_affinity_priority, build_taint_priority]


def main():
    try:
        import golden_more  # noqa: F401  (more transcriptions, same directory)
        BUILDERS.extend(golden_more.builders(globals()))
    except ImportError:
        pass
    try:
        import golden_extra  # noqa: F401  (third batch)
        BUILDERS.extend(golden_extra.builders(globals()))
    except ImportError:
        pass
    try:
        import golden_policy  # noqa: F401  (fourth batch: Policy-enabled priorities and predicate)
        BUILDERS.extend(golden_policy.builders(globals()))
    except ImportError:
        pass
    try:
        import golden_preempt  # noqa: F401  (fifth batch: preemption)
        BUILDERS.extend(golden_preempt.builders(globals()))
    except ImportError:
        pass
    for b in BUILDERS:
        b()
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_vectors.json")
    with open(out, "w") as f:
        json.dump({"cases": CASES}, f, indent=0, sort_keys=True)
    print("wrote %d cases to %s" % (len(CASES), out))


if __name__ == "__main__":
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    main()
