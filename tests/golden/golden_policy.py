"""Fourth batch of transcriptions of the reference's own tests (see build_fixtures.py): the priorities and the
predicate that are registered by name but only enabled through a Policy (algorithmprovider/defaults/defaults.go:93-116,
api/types.go:36-125). Each builder cites its source."""


def builders(g):
    node, pod, ctr = g["node"], g["pod"], g["ctr"]
    CASES = g["CASES"]

    def lift(*keys):  # arguments of the plugin itself belong to the case, not to the cluster
        for k in keys:
            if k in CASES[-1]["cluster"]:
                CASES[-1][k] = CASES[-1]["cluster"].pop(k)

    def priority(*a, **kw):
        g["priority"](*a, **kw)
        lift("args", "policy")

    def predicate(*a, **kw):
        g["predicate"](*a, **kw)
        lift("args", "policy")

    # ----------------------------------------------------------------- MostRequestedPriority
    # algorithm/priorities/most_requested_test.go:30-222
    def build_most_requested():
        src = "algorithm/priorities/most_requested_test.go:107-210"
        l1, l2 = {"foo": "bar", "baz": "blah"}, {"bar": "foo", "baz": "blah"}
        cpu_only = [ctr(1000, 0), ctr(2000, 0)]
        cpu_mem = [ctr(1000, 2000), ctr(2000, 3000)]
        big = [ctr(2000, 4000), ctr(3000, 5000)]
        P = "MostRequestedPriority"
        priority(P, src, "nothing scheduled, nothing requested", pod(containers=[]),
                 [node("machine1", 4000, 10000), node("machine2", 4000, 10000)], [], [("machine1", 0), ("machine2", 0)])
        priority(P, src, "nothing scheduled, resources requested, differently sized machines", pod(containers=cpu_mem),
                 [node("machine1", 4000, 10000), node("machine2", 6000, 10000)], [], [("machine1", 6), ("machine2", 5)])
        priority(P, src, "no resources requested, pods scheduled with resources", pod(containers=[]),
                 [node("machine1", 10000, 20000), node("machine2", 10000, 20000)],
                 [pod("machine1", labels=l2, containers=cpu_only), pod("machine1", labels=l1, containers=cpu_only),
                  pod("machine2", labels=l1, containers=cpu_only), pod("machine2", labels=l1, containers=cpu_mem)],
                 [("machine1", 3), ("machine2", 4)])
        priority(P, src, "resources requested, pods scheduled with resources", pod(containers=cpu_mem),
                 [node("machine1", 10000, 20000), node("machine2", 10000, 20000)],
                 [pod("machine1", containers=cpu_only), pod("machine2", containers=cpu_mem)],
                 [("machine1", 4), ("machine2", 5)])
        priority(P, src, "resources requested with more than the node, pods scheduled with resources", pod(containers=big),
                 [node("machine1", 4000, 10000), node("machine2", 10000, 8000)], [], [("machine1", 4), ("machine2", 2)])

    # ----------------------------------------------------------------- NodeLabelPriority
    # algorithm/priorities/node_label_test.go:30-125 (the pod is nil there: an empty pod here)
    def build_node_label():
        src = "algorithm/priorities/node_label_test.go:42-108"
        nodes = lambda: [node("machine1", labels={"foo": "bar"}), node("machine2", labels={"bar": "foo"}),
                         node("machine3", labels={"bar": "baz"})]
        for label, presence, scores, test in [
                ("baz", True, (0, 0, 0), "no match found, presence true"),
                ("baz", False, (10, 10, 10), "no match found, presence false"),
                ("foo", True, (10, 0, 0), "one match found, presence true"),
                ("foo", False, (0, 10, 10), "one match found, presence false"),
                ("bar", True, (0, 10, 10), "two matches found, presence true"),
                ("bar", False, (10, 0, 0), "two matches found, presence false")]:
            priority("NodeLabelPriority", src, test, pod(containers=[]), nodes(), [],
                     list(zip(("machine1", "machine2", "machine3"), scores)), args={"label": label, "presence": presence})

    # ----------------------------------------------------------------- ResourceLimitsPriority
    # algorithm/priorities/resource_limits_test.go:30-151
    def build_resource_limits():
        src = "algorithm/priorities/resource_limits_test.go:104-137"
        lim = lambda c, m: ctr(limits={"cpu": c, "memory": m})
        cpu_only = [lim(1000, 0), lim(2000, 0)]
        mem_only = [lim(0, 2000), lim(0, 3000)]
        cpu_mem = [lim(1000, 2000), lim(2000, 3000)]
        P = "ResourceLimitsPriority"
        priority(P, src, "pod does not specify its resource limits", pod(containers=[]),
                 [node("machine1", 4000, 10000), node("machine2", 4000, 0), node("machine3", 0, 10000), node("machine4", 0, 0)],
                 [], [("machine1", 0), ("machine2", 0), ("machine3", 0), ("machine4", 0)])
        priority(P, src, "pod only specifies  cpu limits", pod(containers=cpu_only),
                 [node("machine1", 3000, 10000), node("machine2", 2000, 10000)], [], [("machine1", 1), ("machine2", 0)])
        priority(P, src, "pod only specifies  mem limits", pod(containers=mem_only),
                 [node("machine1", 4000, 4000), node("machine2", 5000, 10000)], [], [("machine1", 0), ("machine2", 1)])
        priority(P, src, "pod specifies both cpu and  mem limits", pod(containers=cpu_mem),
                 [node("machine1", 4000, 4000), node("machine2", 5000, 10000)], [], [("machine1", 1), ("machine2", 1)])
        priority(P, src, "node does not advertise its allocatables", pod(containers=cpu_mem), [node("machine1", 0, 0)], [],
                 [("machine1", 0)])

    # ----------------------------------------------------------------- CheckNodeLabelPresence
    # algorithm/predicates/predicates_test.go:1387-1452
    def build_label_presence():
        src = "algorithm/predicates/predicates_test.go:1396-1431"
        for labels, presence, fits, test in [
                (["baz"], True, False, "label does not match, presence true"),
                (["baz"], False, True, "label does not match, presence false"),
                (["foo", "baz"], True, False, "one label matches, presence true"),
                (["foo", "baz"], False, False, "one label matches, presence false"),
                (["foo", "bar"], True, True, "all labels match, presence true"),
                (["foo", "bar"], False, False, "all labels match, presence false")]:
            predicate("CheckNodeLabelPresence", src, test, pod(containers=[]),
                      node("machine1", labels={"foo": "bar", "bar": "foo"}), fits=fits,
                      reasons=None if fits else ["CheckNodeLabelPresence"],
                      policy={"labelsPresence": {"labels": labels, "presence": presence}})

    # ----------------------------------------------------------------- ImageLocalityPriority
    # algorithm/priorities/image_locality_test.go:30-176
    def build_image_locality():
        src = "algorithm/priorities/image_locality_test.go:31-161"
        mb = 1024 * 1024
        img = lambda *names: [{"requests": {}, "image": n} for n in names]
        node_40_140_2000 = [{"names": ["gcr.io/40", "gcr.io/40:v1", "gcr.io/40:v1"], "sizeBytes": 40 * mb},
                            {"names": ["gcr.io/140", "gcr.io/140:v1"], "sizeBytes": 140 * mb},
                            {"names": ["gcr.io/2000"], "sizeBytes": 2000 * mb}]
        node_250_10 = [{"names": ["gcr.io/250"], "sizeBytes": 250 * mb},
                       {"names": ["gcr.io/10", "gcr.io/10:v1"], "sizeBytes": 10 * mb}]
        nodes = lambda: [node("machine1", images=node_40_140_2000), node("machine2", images=node_250_10)]
        P = "ImageLocalityPriority"
        priority(P, src, "two images spread on two nodes, prefer the larger image one",
                 pod(containers=img("gcr.io/40", "gcr.io/250")), nodes(), [], [("machine1", 1), ("machine2", 3)])
        priority(P, src, "two images on one node, prefer this node",
                 pod(containers=img("gcr.io/40", "gcr.io/140")), nodes(), [], [("machine1", 2), ("machine2", 0)])
        priority(P, src, "if exceed limit, use limit",
                 pod(containers=img("gcr.io/10", "gcr.io/2000")), nodes(), [], [("machine1", 10), ("machine2", 0)])

    # ----------------------------------------------------------------- selectHost
    # core/generic_scheduler_test.go:117-181 (TestSelectHost). The HostPriorityList is handed to the whole Schedule path
    # as the combined score of an extender (core/generic_scheduler.go:608-632) with every priority weighted 0, ten pods
    # in a row as the test's ten calls. `host_in` is what the reference asserts; `hosts` is the order selectHost's code
    # implies (:177-192: descending (score, host), lastNodeIndex round-robin over the top scorers, starting at 0).
    # The fourth case of the test (an empty list is an error) cannot occur on this path: an empty filtered list is a
    # FitError before selectHost is reached (:131-137).
    def build_select_host():
        src = "core/generic_scheduler_test.go:117-181"
        zero = {k: 0 for k in ("LeastRequestedPriority", "BalancedResourceAllocation", "SelectorSpreadPriority",
                               "NodeAffinityPriority", "TaintTolerationPriority", "InterPodAffinityPriority",
                               "NodePreferAvoidPodsPriority")}
        for test, lst, possible in [
                ("unique, properly ordered scores", [("machine1.1", 1), ("machine2.1", 2)], ["machine2.1"]),
                ("equal scores", [("machine1.1", 1), ("machine1.2", 2), ("machine1.3", 2), ("machine2.1", 2)],
                 ["machine1.2", "machine1.3", "machine2.1"]),
                ("out of order scores", [("machine1.1", 3), ("machine1.2", 3), ("machine2.1", 2), ("machine3.1", 1),
                                         ("machine1.3", 3)], ["machine1.1", "machine1.2", "machine1.3"])]:
            order = sorted(possible, reverse=True)
            CASES.append({"op": "stream", "source": src, "test": "selectHost: " + test, "weights": zero,
                          "extender": {"scores": dict(lst)},
                          "cluster": {"nodes": [node(h, 4000, 1 << 33, pods=110) for h, _ in lst],
                                      "pods": [pod(containers=[]) for _ in range(10)]},
                          "expect": {"host_in": [possible] * 10, "hosts": [order[i % len(order)] for i in range(10)]}})

    # ----------------------------------------------------------------- GeneralPredicates
    # algorithm/predicates/predicates_test.go:2051-2159 (TestRunGeneralPredicates)
    def build_general_predicates():
        src = "algorithm/predicates/predicates_test.go:2051-2159"
        GPU = "alpha.kubernetes.io/nvidia-gpu"
        rpod = lambda cpu, mem, gpu=None: pod(containers=[ctr(cpu, mem, requests={GPU: gpu} if gpu is not None else None)])
        plain = lambda: node("machine1", alloc={"cpu": 10, "memory": 20, GPU: 0, "pods": 32})
        gpu1 = lambda: node("machine1", alloc={"cpu": 10, "memory": 20, GPU: 1, "pods": 32})
        port_pod = lambda: pod(containers=[{"requests": {}, "ports": [{"hostPort": 123}]}])
        P = "GeneralPredicates"
        predicate(P, src, "no resources/port/host requested always fits", pod(), plain(), [rpod(9, 19)], True, [])
        predicate(P, src, "not enough cpu and memory resource", rpod(8, 10), plain(), [rpod(5, 19)], False,
                  ["Insufficient cpu", "Insufficient memory"])
        predicate(P, src, "no resources/port/host requested always fits on GPU machine", pod(), gpu1(), [rpod(9, 19)], True, [])
        predicate(P, src, "not enough GPU resource", rpod(3, 1, 1), gpu1(), [rpod(5, 10, 1)], False, ["Insufficient " + GPU])
        predicate(P, src, "enough GPU resource", rpod(3, 1, 1), gpu1(), [rpod(5, 10, 0)], True, [])
        predicate(P, src, "host not match", pod(nodeName="machine2"), plain(), [], False, ["HostName"])
        predicate(P, src, "hostport conflict", port_pod(), plain(), [port_pod()], False, ["PodFitsHostPorts"])

    # ----------------------------------------------------------------- FitError text
    # core/generic_scheduler_test.go:410-426 (TestHumanReadableFitError): one node under memory pressure, two under disk
    # pressure; here the FailedPredicateMap comes out of a real Schedule call of a best-effort pod
    def build_fit_error_text():
        cond = lambda *pairs: [{"type": t, "status": s} for t, s in pairs]
        CASES.append({"op": "stream", "source": "core/generic_scheduler_test.go:410-426", "test": "human readable FitError",
                      "cluster": {"nodes": [node("1", 1000, 1 << 30, pods=110, conditions=cond(("MemoryPressure", "True"))),
                                            node("2", 1000, 1 << 30, pods=110, conditions=cond(("DiskPressure", "True"))),
                                            node("3", 1000, 1 << 30, pods=110, conditions=cond(("DiskPressure", "True")))],
                                  "pods": [pod(name="2", containers=[{}])]},
                      "expect": {"status": [2], "message_contains": [["0/3 nodes are available", "2 NodeUnderDiskPressure",
                                                                       "1 NodeUnderMemoryPressure"]]}})

    # ----------------------------------------------------------------- extender verdicts
    # core/extender_test.go:188-337 (TestGenericSchedulerWithExtenders). The fake extenders' verdicts are worked out
    # here as HTTPExtender's caller would hand them on: `keep` = intersection of their Filter results (:108-133),
    # `scores` = sum over extenders of extender weight * sum over its prioritizers of (score * prioritizer weight)
    # (:135-158, generic_scheduler.go:608-632). The built-in side is truePredicate + EqualPriority weight 1, as in the
    # test. "test 1" (an extender returns an error) has no counterpart: the shim fails the call before the engine.
    # In "test 7" the built-in fake machine2Prioritizer (weight 20) is folded into the scores.
    def build_extenders():
        src = "core/extender_test.go:215-306"
        w = {k: 0 for k in ("LeastRequestedPriority", "BalancedResourceAllocation", "SelectorSpreadPriority",
                            "NodeAffinityPriority", "TaintTolerationPriority", "InterPodAffinityPriority",
                            "NodePreferAvoidPodsPriority")}
        w["EqualPriority"] = 1
        m = lambda *names: [node(n, 4000, 1 << 33, pods=110) for n in names]
        for test, names, ext, expect in [
                ("test 2", ("machine1", "machine2"), {"keep": []}, {"status": [2]}),
                ("test 3", ("machine1", "machine2"), {"keep": ["machine1"]}, {"status": [0], "hosts": ["machine1"]}),
                ("test 4", ("machine1", "machine2"), {"keep": []}, {"status": [2]}),
                ("test 5", ("machine1",), {"keep": ["machine1"]}, {"status": [0], "hosts": ["machine1"]}),
                ("test 6", ("machine1", "machine2"),
                 {"keep": ["machine1", "machine2"], "scores": {"machine1": 1 * (10 * 10) + 5 * (1 * 10), "machine2": 1 * (1 * 10) + 5 * (10 * 10)}},
                 {"status": [0], "hosts": ["machine2"], "scores": [511]}),
                ("test 7", ("machine1", "machine2"),
                 {"keep": ["machine1", "machine2"], "scores": {"machine1": 20 * 1 + 1 * (10 * 10), "machine2": 20 * 10 + 1 * (1 * 10)}},
                 {"status": [0], "hosts": ["machine2"]})]:
            CASES.append({"op": "stream", "source": src, "test": "extenders: " + test, "weights": w, "extender": ext,
                          "cluster": {"nodes": m(*names), "pods": [pod(containers=[])]}, "expect": expect})

    # ----------------------------------------------------------------- topology helpers
    # algorithm/priorities/util/topologies_test.go:41-260
    def build_topology_helpers():
        src = "algorithm/priorities/util/topologies_test.go"
        fake = lambda **kw: dict({"name": "topologies_pod", "namespace": "default"}, **kw)

        def helper(test, lines, expect, **req):
            CASES.append(dict({"op": "helper", "source": "%s:%s" % (src, lines), "test": test, "expect": {"result": expect}}, **req))

        helper("podAffinityTerm_namespace_empty", "41-66", ["default"], fn="term_namespaces", pod=fake(), term={})
        helper("podAffinityTerm_namespace_not_empty", "41-66", ["kube-public", "kube-system"], fn="term_namespaces", pod=fake(),
               term={"namespaces": ["kube-public", "kube-system"]})
        sel_ = {"matchExpressions": [{"key": "service", "op": "In", "values": ["topologies_service1", "topologies_service2"]}]}
        for test, ns, labels, want in [("namespace_not_in", "default", {"service": "topologies_service1"}, False),
                                       ("label_not_match", "kube-public", {"service": "topologies_service3"}, False),
                                       ("normal_case", "kube-public", {"service": "topologies_service1"}, True)]:
            helper(test, "68-109", want, fn="ns_selector_match", pod=fake(namespace=ns, labels=labels),
                   namespaces=["kube-public", "kube-system"], selector=sel_)
        n = lambda labels=None: ({"name": "n", "labels": labels} if labels is not None else {"name": "n"})
        for test, a, b, key, want in [
                ("nodeA{'a':'a'} vs. empty label in nodeB", n({"a": "a"}), n(), "a", False),
                ("nodeA{'a':'a'} vs. nodeB{'a':'a'}", n({"a": "a"}), n({"a": "a"}), "a", True),
                ("nodeA{'a':''} vs. empty label in nodeB", n({"a": ""}), n(), "a", False),
                ("nodeA{'a':''} vs. nodeB{'a':''}", n({"a": ""}), n({"a": ""}), "a", True),
                ("nodeA{'a':'a'} vs. nodeB{'a':'a'} by key{'b'}", n({"a": "a"}), n({"a": "a"}), "b", False),
                ("topologyKey empty", n({"a": ""}), n({"a": ""}), "", False),
                ("nodeA lable nil vs. nodeB{'a':''} by key('a')", n(), n({"a": ""}), "a", False),
                ("nodeA{'a':''}  vs. nodeB label is nil by key('a')", n({"a": ""}), n(), "a", False)]:
            helper(test, "111-260", want, fn="same_topology", nodeA=a, nodeB=b, key=key)

    # ----------------------------------------------------------------- small utilities
    # algorithm/priorities/util/non_zero_test.go:27-74, util/utils_test.go:25-104
    def build_utilities():
        def helper(src, test, expect, **req):
            CASES.append(dict({"op": "helper", "source": src, "test": test, "expect": {"result": expect}}, **req))

        DC, DM = 100, 200 * 1024 * 1024
        src = "algorithm/priorities/util/non_zero_test.go:27-74"
        for test, requests, want in [("cpu_and_memory_not_found", {}, [DC, DM]), ("only_cpu_exist", {"cpu": 200}, [200, DM]),
                                     ("only_memory_exist", {"memory": 400 << 20}, [DC, 400 << 20]),
                                     ("cpu_memory_exist", {"cpu": 200, "memory": 400 << 20}, [200, 400 << 20])]:
            helper(src, test, want, fn="nonzero_requests", pod=pod(containers=[{"requests": requests}]))
        src = "util/utils_test.go:25-59"
        helper(src, "no priority pod resolves to static default priority", 0, fn="pod_priority", pod=pod(containers=[{}]))
        helper(src, "pod with priority resolves correctly", 20, fn="pod_priority", pod=pod(containers=[{}], priority=20))
        # algorithm/priorities/metadata_test.go:32-160 (the nil pod of its first case has no counterpart)
        src = "algorithm/priorities/metadata_test.go:118-145"
        tol = [{"key": "foo", "operator": "Equal", "value": "bar", "effect": "PreferNoSchedule"}]
        tol_out = [["foo", "Equal", "bar", "PreferNoSchedule"]]
        aff = {"podAffinity": {"preferred": [{"weight": 5, "term": {"selector": {"matchExpressions": [
            {"key": "security", "op": "In", "values": ["S1"]}]}, "topologyKey": "region"}}]}}
        helper(src, "Produce a priorityMetadata with default requests", {"nonzero": [DC, DM], "tolerations": tol_out},
               fn="priority_meta", pod=pod(containers=[{}], affinity=aff, tolerations=tol))
        helper(src, "Produce a priorityMetadata with specified requests (tolerations)", {"nonzero": [200, 2000], "tolerations": tol_out},
               fn="priority_meta", pod=pod(containers=[ctr(200, 2000)], tolerations=tol))
        helper(src, "Produce a priorityMetadata with specified requests (affinity)", {"nonzero": [200, 2000], "tolerations": []},
               fn="priority_meta", pod=pod(containers=[ctr(200, 2000)], affinity=aff))
        # algorithm/predicates/predicates_test.go:656-689 (TestGetUsedPorts)
        src = "algorithm/predicates/predicates_test.go:656-689"

        def ports_pod(*infos):
            ports = []
            for i in infos:
                proto, ip, port = i.split("/")
                ports.append({"protocol": proto, "hostIP": ip, "hostPort": int(port)})
            return pod(containers=[{"ports": ports}])

        helper(src, "one pod, one port", ["UDP/127.0.0.1/9090"], fn="used_ports", pods=[ports_pod("UDP/127.0.0.1/9090")])
        helper(src, "two pods on one machine", ["UDP/127.0.0.1/9090", "UDP/127.0.0.1/9091"], fn="used_ports",
               pods=[ports_pod("UDP/127.0.0.1/9090"), ports_pod("UDP/127.0.0.1/9091")])
        helper(src, "two pods, two protocols", ["TCP/0.0.0.0/9090", "UDP/127.0.0.1/9091"], fn="used_ports",
               pods=[ports_pod("TCP/0.0.0.0/9090"), ports_pod("UDP/127.0.0.1/9091")])
        # staging/src/k8s.io/api/core/v1/toleration_test.go:23-125 (TestTolerationToleratesTaint); paths outside the
        # scheduler tree are given from the repository root
        src = "/staging/src/k8s.io/api/core/v1/toleration_test.go:23-125"
        NS, NE = "NoSchedule", "NoExecute"
        for test, tol_, taint, want in [
                ("same key and effect, Exists, taint has no value", ("foo", "Exists", "", NS), ("foo", "", NS), True),
                ("same key and effect, Exists, taint has some value", ("foo", "Exists", "", NS), ("foo", "bar", NS), True),
                ("empty key with Exists matches all taints", ("", "Exists", "", NS), ("foo", "bar", NS), True),
                ("same key, effect and value, Equal", ("foo", "Equal", "bar", NS), ("foo", "bar", NS), True),
                ("different values, Equal", ("foo", "Equal", "value1", NS), ("foo", "value2", NS), False),
                ("different effects, Equal", ("foo", "Equal", "bar", NS), ("foo", "bar", NE), False)]:
            helper(src, test, want, fn="tolerates_taint",
                   toleration=dict(zip(("key", "operator", "value", "effect"), tol_)), taint=dict(zip(("key", "value", "effect"), taint)))
        # staging/src/k8s.io/apimachinery/pkg/labels/selector_test.go:379-425 (TestRequirementSelectorMatching; its first
        # case, a zero-value Requirement, cannot be built through the constructors the path uses)
        src = "/staging/src/k8s.io/apimachinery/pkg/labels/selector_test.go:379-425"
        rq = lambda key, op, *values: {"key": key, "op": op, "values": list(values)}
        for k, (labels_, reqs_, want) in enumerate([
                ({"x": "foo", "y": "baz"}, [rq("x", "In", "foo"), rq("y", "NotIn", "alpha")], True),
                ({"x": "foo", "y": "baz"}, [rq("x", "In", "foo"), rq("y", "In", "alpha")], False),
                ({"y": ""}, [rq("x", "NotIn", ""), rq("y", "Exists")], True),
                ({"y": ""}, [rq("x", "DoesNotExist"), rq("y", "Exists")], True),
                ({"y": ""}, [rq("x", "NotIn", ""), rq("y", "DoesNotExist")], False),
                ({"y": "baz"}, [rq("x", "In", "")], False),
                ({"z": "2"}, [rq("z", "Gt", "1")], True),
                ({"z": "v2"}, [rq("z", "Gt", "1")], False)]):
            helper(src, "requirement selector matching, case %d" % (k + 2), want, fn="requirements_match", labels=labels_, reqs=reqs_)
        # pkg/apis/core/v1/helper/helpers_test.go:246-308 (TestNodeSelectorRequirementsAsSelector)
        src = "/pkg/apis/core/v1/helper/helpers_test.go:246-308"
        for test, reqs_, want in [
                ("no requirements select nothing", [], {"nothing": True, "reqs": []}),
                ("In with two values", [rq("foo", "In", "bar", "baz")], {"nothing": False, "reqs": [["foo", "In", ["bar", "baz"]]]}),
                ("Exists with values is an error", [rq("foo", "Exists", "bar", "baz")], {"error": True}),
                ("Gt", [rq("foo", "Gt", "1")], {"nothing": False, "reqs": [["foo", "Gt", ["1"]]]}),
                ("Lt", [rq("bar", "Lt", "7")], {"nothing": False, "reqs": [["bar", "Lt", ["7"]]]})]:
            helper(src, "NodeSelectorRequirementsAsSelector: " + test, want, fn="requirements_selector", reqs=reqs_)
        # staging/src/k8s.io/apimachinery/pkg/apis/meta/v1/helpers_test.go:27-84 (TestLabelSelectorAsSelector)
        src = "/staging/src/k8s.io/apimachinery/pkg/apis/meta/v1/helpers_test.go:27-84"
        ml, me = {"foo": "bar"}, [rq("baz", "In", "qux", "norf")]
        for test, sel_in, want in [
                ("nil selects nothing", None, {"nothing": True, "reqs": []}),
                ("empty selects everything", {}, {"nothing": False, "reqs": []}),
                ("matchLabels", {"matchLabels": ml}, {"nothing": False, "reqs": [["foo", "Equals", ["bar"]]]}),
                ("matchExpressions", {"matchExpressions": me}, {"nothing": False, "reqs": [["baz", "In", ["norf", "qux"]]]}),
                ("both", {"matchLabels": ml, "matchExpressions": me},
                 {"nothing": False, "reqs": [["baz", "In", ["norf", "qux"]], ["foo", "Equals", ["bar"]]]}),
                ("Exists with values is an error", {"matchExpressions": [rq("baz", "Exists", "qux", "norf")]}, {"error": True})]:
            helper(src, "LabelSelectorAsSelector: " + test, want, fn="label_selector", selector=sel_in)
        # TestSortableList: ten pods of priorities 0..9 in that order, sorted by HigherPriorityPod: 9 first
        helper("util/utils_test.go:61-104", "SortableList sorts by decreasing priority", list(range(9, -1, -1)),
               fn="sort_by_priority", priorities=list(range(10)))

    return [build_most_requested, build_node_label, build_resource_limits, build_label_presence, build_image_locality,
            build_select_host, build_general_predicates, build_fit_error_text, build_extenders, build_topology_helpers,
            build_utilities]
