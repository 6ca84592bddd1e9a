"""Third batch of transcriptions of the reference's own tests (see build_fixtures.py): the small predicate and
priority tables, and the cache aggregates. Each builder cites its source."""


def builders(g):
    node, pod, ctr, expr = g["node"], g["pod"], g["ctr"], g["expr"]
    priority, predicate, CASES = g["priority"], g["predicate"], g["CASES"]

    # ----------------------------------------------------------------- NodePreferAvoidPodsPriority
    # algorithm/priorities/node_prefer_avoid_pods_test.go:30-145
    def build_prefer_avoid():
        src = "algorithm/priorities/node_prefer_avoid_pods_test.go:84-128"
        nodes = lambda: [node("machine1", avoid=[{"kind": "ReplicationController", "uid": "abcdef123456"}]),
                         node("machine2", avoid=[{"kind": "ReplicaSet", "uid": "qwert12345"}]),
                         node("machine3")]
        P = "NodePreferAvoidPodsPriority"
        priority(P, src, "pod managed by ReplicationController should avoid a node, this node get lowest priority score",
                 pod(ns="default", controller={"kind": "ReplicationController", "uid": "abcdef123456"}), nodes(), [],
                 [("machine1", 0), ("machine2", 10), ("machine3", 10)])
        priority(P, src, "ownership by random controller should be ignored",
                 pod(ns="default", controller={"kind": "RandomController", "uid": "abcdef123456"}), nodes(), [],
                 [("machine1", 10), ("machine2", 10), ("machine3", 10)])
        # an owner reference without Controller=true is not a controller reference: the pod has none
        priority(P, src, "owner without Controller field set should be ignored", pod(ns="default"), nodes(), [],
                 [("machine1", 10), ("machine2", 10), ("machine3", 10)])
        priority(P, src, "pod managed by ReplicaSet should avoid a node, this node get lowest priority score",
                 pod(ns="default", controller={"kind": "ReplicaSet", "uid": "qwert12345"}), nodes(), [],
                 [("machine1", 10), ("machine2", 0), ("machine3", 10)])

    # ----------------------------------------------------------------- hard pod affinity symmetric weight
    # algorithm/priorities/interpod_affinity_test.go:529-612
    def build_hard_affinity_weight():
        src = "algorithm/priorities/interpod_affinity_test.go:566-594"
        hard = {"podAffinity": {"required": [{"selector": {"matchExpressions": [expr("service", "In", "S1")]},
                                              "topologyKey": "region"}]}}
        nodes = lambda: [node("machine1", labels={"region": "China"}), node("machine2", labels={"region": "India"}),
                         node("machine3", labels={"az": "az1"})]
        existing = lambda: [pod("machine1", affinity=hard), pod("machine2", affinity=hard)]
        for weight, scores, test in [
                (1, [("machine1", 10), ("machine2", 10), ("machine3", 0)],
                 "Hard Pod Affinity symmetry: weight 1 by default, nodes matching the symmetry rules get a high score"),
                (0, [("machine1", 0), ("machine2", 0), ("machine3", 0)],
                 "Hard Pod Affinity symmetry: closed (weight 0), matching nodes score like the others")]:
            priority("InterPodAffinityPriority", src, test, pod(labels={"service": "S1"}), nodes(), existing(), scores,
                     )
            CASES[-1]["hardPodAffinityWeight"] = weight

    # ----------------------------------------------------------------- PodFitsHost
    # algorithm/predicates/predicates_test.go:458-516
    def build_fits_host():
        src = "algorithm/predicates/predicates_test.go:465-496"
        predicate("PodFitsHost", src, "no host specified", pod(), node(""), fits=True)
        predicate("PodFitsHost", src, "host matches", pod(nodeName="foo"), node("foo"), fits=True)
        predicate("PodFitsHost", src, "host doesn't match", pod(nodeName="bar"), node("foo"), fits=False,
                  reasons=["HostName"])

    # ----------------------------------------------------------------- memory / disk pressure, node conditions
    # algorithm/predicates/predicates_test.go:3422-3686
    def build_pressure_and_conditions():
        cond = lambda *pairs: [{"type": t, "status": s} for t, s in pairs]
        src = "algorithm/predicates/predicates_test.go:3481-3506"
        best_effort = lambda: pod(containers=[{}])
        burstable = lambda: pod(containers=[ctr(100, 100, requests={"alpha.kubernetes.io/nvidia-gpu": 100, "pods": 100})])
        no_mp = lambda: node("n", conditions=cond(("Ready", "True")))
        mp = lambda: node("n", conditions=cond(("MemoryPressure", "True")))
        M = "CheckNodeMemoryPressure"
        predicate(M, src, "best-effort pod schedulable on node without memory pressure condition on", best_effort(), no_mp())
        predicate(M, src, "best-effort pod not schedulable on node with memory pressure condition on", best_effort(), mp(),
                  fits=False, reasons=["NodeUnderMemoryPressure"])
        predicate(M, src, "non best-effort pod schedulable on node with memory pressure condition on", burstable(), mp())
        predicate(M, src, "non best-effort pod schedulable on node without memory pressure condition on", burstable(), no_mp())
        src = "algorithm/predicates/predicates_test.go:3565-3580"
        D = "CheckNodeDiskPressure"
        predicate(D, src, "pod schedulable on node without pressure condition on", pod(containers=[{}]),
                  node("n", conditions=cond(("Ready", "True"))))
        predicate(D, src, "pod not schedulable on node with pressure condition on", pod(containers=[{}]),
                  node("n", conditions=cond(("DiskPressure", "True"))), fits=False, reasons=["NodeUnderDiskPressure"])
        # TestNodeConditionPredicate pins fit only; the reason texts follow predicates.go:1414-1441 and error.go:44-48
        src = "algorithm/predicates/predicates_test.go:3598-3660"
        C = "CheckNodeCondition"
        table = [
            ("node1", cond(("Ready", "True")), False, []),
            ("node2", cond(("Ready", "False")), False, ["NodeNotReady"]),
            ("node3", cond(("OutOfDisk", "True")), False, ["NodeOutOfDisk"]),
            ("node4", cond(("OutOfDisk", "False")), False, []),
            ("node5", cond(("Ready", "True"), ("OutOfDisk", "True")), False, ["NodeOutOfDisk"]),
            ("node6", cond(("Ready", "True"), ("OutOfDisk", "False")), False, []),
            ("node7", cond(("Ready", "False"), ("OutOfDisk", "True")), False, ["NodeNotReady", "NodeOutOfDisk"]),
            ("node8", cond(("Ready", "False"), ("OutOfDisk", "False")), False, ["NodeNotReady"]),
            ("node9", [], True, ["NodeUnschedulable"]),
            ("node10", [], False, []),
            ("node11", [], False, []),
        ]
        for name, conditions, unsched, reasons in table:
            n = node(name, conditions=conditions)
            if unsched:
                n["unschedulable"] = True
            predicate(C, src, name, pod(), n, fits=not reasons, reasons=sorted(reasons))

    # ----------------------------------------------------------------- cache aggregates
    # schedulercache/cache_test.go:48-189 (TestAssumePodScheduled): NodeInfo after AssumePod of the listed pods
    def build_assume_pod():
        src = "schedulercache/cache_test.go:48-189"
        port = lambda p: {"hostIP": "127.0.0.1", "hostPort": p, "protocol": "TCP"}

        def base(name, cpu, mem, ext, ports):
            req = {}
            if cpu is not None:
                req["cpu"] = cpu
            if mem is not None:
                req["memory"] = mem
            if ext:
                req[ext[0]] = ext[1]
            return pod("node", ns="node-namespace", name=name, containers=[ctr(requests=req, ports=ports)])

        tp = [base("test", 100, 500, None, [port(80)]), base("test-1", 100, 500, None, [port(80)]),
              base("test-2", 200, 1024, None, [port(8080)]), base("test-nonzero", None, None, None, [port(80)]),
              base("test", 100, 500, ("example.com/foo", 3), [port(80)]),
              base("test-2", 200, 1024, ("example.com/foo", 5), [port(8080)]),
              base("test", 100, 500, ("random-invalid-extended-key", 100), [{}])]
        table = [
            ([0], (100, 500, {}), (100, 500), ["TCP/127.0.0.1/80"]),
            ([1, 2], (300, 1524, {}), (300, 1524), ["TCP/127.0.0.1/80", "TCP/127.0.0.1/8080"]),
            ([3], (0, 0, {}), (100, 200 * 1024 * 1024), ["TCP/127.0.0.1/80"]),
            ([4], (100, 500, {"example.com/foo": 3}), (100, 500), ["TCP/127.0.0.1/80"]),
            ([4, 5], (300, 1524, {"example.com/foo": 8}), (300, 1524), ["TCP/127.0.0.1/80", "TCP/127.0.0.1/8080"]),
            ([6], (100, 500, {}), (100, 500), []),
        ]
        for i, (idx, req, nz, ports) in enumerate(table):
            CASES.append({"op": "node_info", "source": src, "test": "TestAssumePodScheduled #%d" % i, "node": "node",
                          "cluster": {"nodes": [], "existingPods": [tp[k] for k in idx]}, "forget": [],
                          "expect": {"requested": {"cpu": req[0], "memory": req[1], "gpu": 0, "ephemeral": 0, "scalars": req[2]},
                                     "nonzero": {"cpu": nz[0], "memory": nz[1]}, "pods": len(idx), "usedPorts": ports}})

    # ----------------------------------------------------------------- cache state machine
    # schedulercache/cache_test.go:200-690: Assume / FinishBinding / Add / Update / Remove / Forget / expiry.
    # Times are seconds from the test's `now`; ttl = 10 s as in the tests.
    def build_cache_ops():
        port = lambda p, ip="127.0.0.1": {"hostIP": ip, "hostPort": p, "protocol": "TCP"}

        def base(node_, name, cpu, mem, ports):
            return pod(node_, ns="node-namespace", name=name, containers=[ctr(requests={"cpu": cpu, "memory": mem}, ports=ports)])

        def info(cpu, mem, npods, ports):
            return {"requested": {"cpu": cpu, "memory": mem, "gpu": 0, "ephemeral": 0, "scalars": {}},
                    "nonzero": {"cpu": cpu, "memory": mem}, "pods": npods, "usedPorts": ports,
                    "allocatable": {"cpu": 0, "memory": 0, "pods": 0, "scalars": {}}, "available": {}}

        def case(src, test, ops, nodes, ttl=10):
            CASES.append({"op": "cache_ops", "source": src, "test": test, "ttl": ttl, "ops": ops, "report": sorted(nodes),
                          "expect": {"errors": [""] * len(ops), "nodes": nodes}})

        assume = lambda p, t: [{"do": "assume", "pod": p}, {"do": "finish", "pod": p, "now": t}]
        t1 = base("node", "test-1", 100, 500, [port(80)])
        t2 = base("node", "test-2", 200, 1024, [port(8080)])
        src = "schedulercache/cache_test.go:200-256"
        case(src, "TestExpirePod #0: assumed pod would expire", assume(t1, 0) + [{"do": "cleanup", "now": 20}], {"node": None})
        case(src, "TestExpirePod #1: first one would expire, second one would not",
             assume(t1, 0) + assume(t2, 15) + [{"do": "cleanup", "now": 20}], {"node": info(200, 1024, 1, ["TCP/127.0.0.1/8080"])})
        src = "schedulercache/cache_test.go:258-309"
        case(src, "TestAddPodWillConfirm: the Add()ed pod is confirmed, the other assumed pod expires",
             assume(t1, 0) + assume(t2, 0) + [{"do": "add", "pod": t1}, {"do": "cleanup", "now": 20}],
             {"node": info(100, 500, 1, ["TCP/127.0.0.1/80"])})
        src = "schedulercache/cache_test.go:311-372"
        wild = lambda p: {"hostPort": p}
        assumed = base("assumed-node-1", "test-1", 100, 500, [wild(80)])
        added = base("actual-node", "test-1", 100, 500, [wild(80)])
        updated = base("actual-node", "test-1", 200, 500, [wild(90)])
        case(src, "TestAddPodWillReplaceAssumed: added to another node than assumed, then updated",
             assume(assumed, 0) + [{"do": "add", "pod": added}, {"do": "update", "old": added, "pod": updated}],
             {"assumed-node-1": None, "actual-node": info(200, 500, 1, ["TCP/0.0.0.0/90"])})
        src = "schedulercache/cache_test.go:374-420"
        t = base("node", "test", 100, 500, [port(80)])
        u = base("node", "test", 200, 1024, [port(8080)])
        case(src, "TestAddPodAfterExpiration: expired, then added back",
             assume(t, 0) + [{"do": "cleanup", "now": 20}, {"do": "add", "pod": t}], {"node": info(100, 500, 1, ["TCP/127.0.0.1/80"])})
        src = "schedulercache/cache_test.go:422-488"
        case(src, "TestUpdatePod: added, updated once", [{"do": "add", "pod": t}, {"do": "update", "old": t, "pod": u}],
             {"node": info(200, 1024, 1, ["TCP/127.0.0.1/8080"])})
        case(src, "TestUpdatePod: added, updated twice",
             [{"do": "add", "pod": t}, {"do": "update", "old": t, "pod": u}, {"do": "update", "old": u, "pod": t}],
             {"node": info(100, 500, 1, ["TCP/127.0.0.1/80"])})
        src = "schedulercache/cache_test.go:490-565"
        case(src, "TestExpireAddUpdatePod: assumed, expired, added, updated twice",
             assume(t, 0) + [{"do": "cleanup", "now": 20}, {"do": "add", "pod": t}, {"do": "update", "old": t, "pod": u},
                             {"do": "update", "old": u, "pod": t}], {"node": info(100, 500, 1, ["TCP/127.0.0.1/80"])})
        src = "schedulercache/cache_test.go:567-608"
        case(src, "TestRemovePod: after Add", [{"do": "add", "pod": t}], {"node": info(100, 500, 1, ["TCP/127.0.0.1/80"])}, ttl=1)
        case(src, "TestRemovePod: after Add and Remove", [{"do": "add", "pod": t}, {"do": "remove", "pod": t}], {"node": None}, ttl=1)
        src = "schedulercache/cache_test.go:610-670"
        ops = assume(t, 0) + [{"do": "is_assumed", "pod": t}, {"do": "forget", "pod": t}, {"do": "is_assumed", "pod": t},
                              {"do": "cleanup", "now": 20}]
        CASES.append({"op": "cache_ops", "source": src, "test": "TestForgetPod", "ttl": 10, "ops": ops, "report": ["node"],
                      "expect": {"errors": ["", "", "assumed", "", "not assumed", ""], "nodes": {"node": None}}})
        # error returns of the state machine (cache.go:127-129, :176-178, :188-190, :252-254, :281-283, :308-310)
        src = "schedulercache/cache.go:118-312"
        k = "node-namespace/test"
        other = base("elsewhere", "test", 100, 500, [port(80)])
        ops = [{"do": "assume", "pod": t}, {"do": "assume", "pod": t}, {"do": "update", "old": t, "pod": u},
               {"do": "remove", "pod": t}, {"do": "forget", "pod": other}, {"do": "add", "pod": t}, {"do": "add", "pod": t},
               {"do": "forget", "pod": t}]
        CASES.append({"op": "cache_ops", "source": src, "test": "error texts of invalid transitions", "ttl": 10, "ops": ops,
                      "report": ["node"],
                      "expect": {"errors": ["", "pod %s state wasn't initial but get assumed" % k,
                                            "pod %s state wasn't added but get updated" % k,
                                            "pod state wasn't added but get removed. Pod key: %s" % k,
                                            "pod %s state was assumed on a different node" % k, "",
                                            "pod was already in added state. Pod key: %s" % k,
                                            "pod %s state wasn't assumed but get forgotten" % k],
                                 "nodes": {"node": info(100, 500, 1, ["TCP/127.0.0.1/80"])}}})

    # ----------------------------------------------------------------- fork: device bookkeeping of the cache
    # schedulercache/node_info_test.go:111-186 (TestAddAndRemovePodExtendedResources), driven through the cache
    # (the Go test calls NodeInfo.AddPod directly; pods get names here because the cache keys them)
    def build_device_cache():
        src = "schedulercache/node_info_test.go:111-186"
        r1, r2 = "nvidia.com/gpu", "nvidia.com/bar"
        ids = ["someId%d" % i for i in range(32)]
        devs = lambda: [{"id": i, "health": "Healthy", "attributes": {}} for i in ids]
        n = node("n", extendedResources={r1: devs(), r2: devs()})
        pods_ = []
        for i in range(16):
            pods_.append(pod("n", ns="d", name="p%d" % i, containers=[{}],
                             extendedResources=[{"name": "foo", "limits": {r1: 1}, "assigned": ["someId%d" % i]}]))
        for i in range(16, 24):
            pods_.append(pod("n", ns="d", name="p%d" % i, containers=[{}],
                             extendedResources=[{"name": "foo", "limits": {r1: 2}, "assigned": ["someId%d" % i, "someId%d" % (i + 8)]}]))
        adds = [{"do": "add_node", "node": n}] + [{"do": "add", "pod": p} for p in pods_]
        removes = [{"do": "remove", "pod": p} for p in pods_]
        mib200 = 200 * 1024 * 1024

        def info(npods, avail1):
            return {"requested": {"cpu": 0, "memory": 0, "gpu": 0, "ephemeral": 0, "scalars": {}},
                    "nonzero": {"cpu": 100 * npods, "memory": mib200 * npods}, "pods": npods, "usedPorts": [],
                    "allocatable": {"cpu": 0, "memory": 0, "pods": 0, "scalars": {}},
                    "available": {r2: sorted(ids), r1: sorted(avail1)}}

        for test, ops, expect in [("after adding 24 pods every nvidia.com/gpu device is used", adds, info(24, [])),
                                  ("after removing them all 32 devices are available again", adds + removes, info(0, ids))]:
            CASES.append({"op": "cache_ops", "source": src, "test": "TestAddAndRemovePodExtendedResources: " + test, "ttl": 30,
                          "ops": ops, "report": ["n"], "expect": {"errors": [""] * len(ops), "nodes": {"n": expect}}})

    # ----------------------------------------------------------------- fork: SetNode over a changing device set
    # schedulercache/node_info_test.go:36-109 (TestSetNode), driven through the cache's AddNode / UpdateNode
    def build_set_node():
        src = "schedulercache/node_info_test.go:36-109"
        r1, r2 = "nvidia.com/gpu", "nvidia.com/bar"
        dev = lambda i: {"id": i, "health": "Healthy", "attributes": {}}
        v1_ = node("n", extendedResources={r1: [dev("someId1"), dev("someId2")]})
        v2_ = node("n", extendedResources={r1: [dev("someId2"), dev("someId3")], r2: [dev("someId1")]})
        v3_ = node("n", extendedResources={r2: [dev("someId1")]})
        ops = [{"do": "add_node", "node": v1_}, {"do": "update_node", "node": v2_}, {"do": "update_node", "node": v3_}]

        def info(avail):
            return {"requested": {"cpu": 0, "memory": 0, "gpu": 0, "ephemeral": 0, "scalars": {}},
                    "nonzero": {"cpu": 0, "memory": 0}, "pods": 0, "usedPorts": [],
                    "allocatable": {"cpu": 0, "memory": 0, "pods": 0, "scalars": {}}, "available": avail}

        for k, (test, avail) in enumerate([("two devices of one resource", {r1: ["someId1", "someId2"]}),
                                           ("a device replaced, a second resource appears", {r1: ["someId2", "someId3"], r2: ["someId1"]}),
                                           ("a resource the node no longer reports keeps an empty domain", {r1: [], r2: ["someId1"]})]):
            CASES.append({"op": "cache_ops", "source": src, "test": "TestSetNode: " + test, "ttl": 30, "ops": ops[:k + 1],
                          "report": ["n"], "expect": {"errors": [""] * (k + 1), "nodes": {"n": info(avail)}}})

    # ----------------------------------------------------------------- node add / update / remove
    # schedulercache/cache_test.go:700-893 (TestNodeOperators). Quantities as the API reads them: memory "100m" and
    # "50m" are fractions of a byte and round up to 1 (Quantity.Value()).
    def build_node_operators():
        src = "schedulercache/cache_test.go:700-893"
        taint = [{"key": "test-key", "value": "test-value", "effect": "PreferNoSchedule"}]
        mk_node = lambda mem: node("test-node", alloc={"cpu": 1000, "memory": mem, "example.com/foo": 1}, taints=taint)
        with_port = pod("test-node", name="pod1", containers=[ctr(500, 1, ports=[{"hostPort": 80}])])
        plain = lambda name: pod("test-node", name=name, containers=[ctr(500, 1)])

        def info(npods, ports, alloc):
            return {"requested": {"cpu": 500 * npods, "memory": npods, "gpu": 0, "ephemeral": 0, "scalars": {}},
                    "nonzero": {"cpu": 500 * npods, "memory": npods}, "pods": npods, "usedPorts": ports,
                    "allocatable": alloc, "available": {}}

        full = {"cpu": 1000, "memory": 1, "pods": 0, "scalars": {"example.com/foo": 1}}
        none = {"cpu": 0, "memory": 0, "pods": 0, "scalars": {}}
        for label, pods_, ports in [("one pod with a host port", [with_port], ["TCP/0.0.0.0/80"]),
                                    ("two pods", [plain("pod1"), plain("pod2")], [])]:
            adds = [{"do": "add_node", "node": mk_node(1)}] + [{"do": "add", "pod": p} for p in pods_]
            for step, ops, alloc in [("case 1: node and pods added", adds, full),
                                     ("case 3: node updated", adds + [{"do": "update_node", "node": mk_node(1)}], full),
                                     ("case 4: a removed node keeps its NodeInfo while pods remain",
                                      adds + [{"do": "remove_node", "name": "test-node"}], none)]:
                CASES.append({"op": "cache_ops", "source": src, "test": "TestNodeOperators (%s) %s" % (label, step), "ttl": 1,
                              "ops": ops, "report": ["test-node"],
                              "expect": {"errors": [""] * len(ops), "nodes": {"test-node": info(len(pods_), ports, alloc)}}})

    return [build_prefer_avoid, build_hard_affinity_weight, build_fits_host, build_pressure_and_conditions, build_assume_pod,
            build_cache_ops, build_device_cache, build_set_node, build_node_operators]
